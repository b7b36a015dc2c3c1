/*
 * oracle/gen.c -- CPU ORACLE (test infrastructure, NOT the product path).
 * See gen.h for the reference lines each function restates.
 */
#include "gen.h"

#include <pthread.h>
#include <stdlib.h>
#include <string.h>

static inline uint32_t getbit(const uint64_t *x, uint32_t i) { return (uint32_t)((x[i >> 6] >> (i & 63)) & 1); }

/* rng.py:38-40: x[idx] <= XOR of x[i] for i in taps(idx), all rows from the old state. */
void orc_lutopt_step(uint32_t n, const uint32_t *row_ptr, const uint16_t *taps,
                     const uint64_t *x_in, uint64_t *x_out)
{
    uint32_t nw = (n + 63) / 64;
    uint64_t tmp[8] = { 0 };
    for (uint32_t r = 0; r < n; r++) {
        uint32_t b = 0;
        for (uint32_t j = row_ptr[r]; j < row_ptr[r + 1]; j++) b ^= getbit(x_in, taps[j]);
        tmp[r >> 6] |= (uint64_t)b << (r & 63);
    }
    memcpy(x_out, tmp, nw * sizeof(uint64_t));
}

/* ---- dense n x n GF(2) matrices, row-major rows of 8 words (n <= 512) ---- */
typedef struct { uint64_t r[512][8]; } dmat;

static void dmat_from_taps(dmat *m, uint32_t n, const uint32_t *row_ptr, const uint16_t *taps)
{
    memset(m, 0, sizeof *m);
    for (uint32_t r = 0; r < n; r++)
        for (uint32_t j = row_ptr[r]; j < row_ptr[r + 1]; j++)
            m->r[r][taps[j] >> 6] ^= (uint64_t)1 << (taps[j] & 63); /* XOR: a repeated tap cancels */
}

/* c = a.b : row r of c = XOR over set bits j of a[r] of row j of b */
static void dmat_mul(dmat *c, const dmat *a, const dmat *b, uint32_t n)
{
    uint32_t nw = (n + 63) / 64;
    for (uint32_t r = 0; r < n; r++) {
        uint64_t acc[8] = { 0 };
        for (uint32_t j = 0; j < n; j++)
            if ((a->r[r][j >> 6] >> (j & 63)) & 1)
                for (uint32_t w = 0; w < nw; w++) acc[w] ^= b->r[j][w];
        memcpy(c->r[r], acc, sizeof acc);
    }
}

static void dmat_apply(const dmat *a, uint32_t n, const uint64_t *x, uint64_t *y)
{
    uint32_t nw = (n + 63) / 64;
    uint64_t out[8] = { 0 };
    for (uint32_t r = 0; r < n; r++) {
        uint64_t p = 0;
        for (uint32_t w = 0; w < nw; w++) p ^= a->r[r][w] & x[w];
        out[r >> 6] |= (uint64_t)(__builtin_popcountll(p) & 1) << (r & 63);
    }
    memcpy(y, out, nw * sizeof(uint64_t));
}

void orc_lutopt_jump(uint32_t n, const uint32_t *row_ptr, const uint16_t *taps,
                     uint64_t e, const uint64_t *x_in, uint64_t *x_out)
{
    uint32_t nw = (n + 63) / 64;
    dmat *sq = (dmat *)malloc(sizeof(dmat)), *tmp = (dmat *)malloc(sizeof(dmat));
    uint64_t x[8] = { 0 };
    memcpy(x, x_in, nw * sizeof(uint64_t));
    dmat_from_taps(sq, n, row_ptr, taps);
    while (e) {
        if (e & 1) dmat_apply(sq, n, x, x);
        e >>= 1;
        if (e) {
            dmat_mul(tmp, sq, sq, n);
            dmat *t = sq; sq = tmp; tmp = t;
        }
    }
    memcpy(x_out, x, nw * sizeof(uint64_t));
    free(sq);
    free(tmp);
}

/* rng.py:96-105 / :173-181: log2(n) levels of y[j] = x[2j] - x[2j+1]. */
int32_t orc_clt(uint32_t n, const uint64_t *x, int model)
{
    int16_t lvl[512];
    uint32_t logn = 0;
    while ((1u << logn) < n) logn++;
    for (uint32_t i = 0; i < n; i++) lvl[i] = (int16_t)getbit(x, i);
    /* CLTGRNG needs n a power of two (rng.py:73-74).  A caller that only wants the LUTOPT words of another n (192)
     * still passes through here: pad with zeros so that the tree reads defined values and |g| <= n/2 holds
     * (uninitialised padding used to index the histogram out of bounds). */
    for (uint32_t i = n; i < (1u << logn); i++) lvl[i] = 0;
    for (uint32_t level = 0; level < logn; level++) {
        uint32_t level_n = 1u << (logn - level);
        for (uint32_t pair = 0; pair < level_n; pair += 2) lvl[pair / 2] = (int16_t)(lvl[pair] - lvl[pair + 1]);
    }
    int32_t g = lvl[0];
    if (model == ORC_MODEL_GATEWARE) {
        /* Signal((logn, True)): keep logn bits, two's complement (rng.py:78,108) */
        int32_t half = (int32_t)(n / 2);
        g = ((g + half) & (int32_t)(n - 1)) - half;
    }
    return g;
}

void orc_clt_run(uint32_t n, const uint32_t *row_ptr, const uint16_t *taps,
                 uint64_t *state, uint64_t count, int model, uint32_t noise_var,
                 uint64_t *hist, int64_t *sum, uint64_t *sumsq,
                 int8_t *out8, int16_t *out16, uint64_t *out_words)
{
    uint32_t nw = (n + 63) / 64;
    int32_t half = (int32_t)(n / 2);
    for (uint64_t i = 0; i < count; i++) {
        orc_lutopt_step(n, row_ptr, taps, state, state);
        int32_t g = orc_clt(n, state, model);
        if (hist) hist[g + half] += 1;
        if (sum) *sum += g;
        if (sumsq) *sumsq += (uint64_t)((int64_t)g * g);
        if (out8) out8[i] = (int8_t)g;
        if (out16) out16[i] = (int16_t)(g * (int32_t)noise_var);
        if (out_words) memcpy(out_words + i * nw, state, nw * sizeof(uint64_t));
    }
}

typedef struct {
    uint32_t n; const uint32_t *row_ptr; const uint16_t *taps;
    uint64_t x[8]; uint64_t skip, count; int model;
    uint64_t hist[513]; int64_t sum; uint64_t sumsq;
} clt_job;

static void *clt_worker(void *arg)
{
    clt_job *j = (clt_job *)arg;
    orc_lutopt_jump(j->n, j->row_ptr, j->taps, j->skip, j->x, j->x);
    orc_clt_run(j->n, j->row_ptr, j->taps, j->x, j->count, j->model, 1, j->hist, &j->sum, &j->sumsq,
                NULL, NULL, NULL);
    return NULL;
}

void orc_clt_run_mt(uint32_t n, const uint32_t *row_ptr, const uint16_t *taps,
                    const uint64_t *x0, uint64_t first, uint64_t count, int model,
                    uint64_t *hist, int64_t *sum, uint64_t *sumsq, int threads)
{
    if (threads < 1) threads = 1;
    uint32_t nw = (n + 63) / 64;
    clt_job *jobs = (clt_job *)calloc((size_t)threads, sizeof(clt_job));
    pthread_t *tid = (pthread_t *)calloc((size_t)threads, sizeof(pthread_t));
    uint64_t per = count / (uint64_t)threads, extra = count % (uint64_t)threads, at = first;
    for (int t = 0; t < threads; t++) {
        clt_job *j = &jobs[t];
        j->n = n; j->row_ptr = row_ptr; j->taps = taps; j->model = model;
        memcpy(j->x, x0, nw * sizeof(uint64_t));
        j->skip = at;
        j->count = per + ((uint64_t)t < extra ? 1 : 0);
        at += j->count;
        pthread_create(&tid[t], NULL, clt_worker, j);
    }
    uint32_t bins = n + (model == ORC_MODEL_NUMPY ? 1 : 0);
    for (int t = 0; t < threads; t++) {
        pthread_join(tid[t], NULL);
        if (hist) for (uint32_t b = 0; b < bins; b++) hist[b] += jobs[t].hist[b];
        if (sum) *sum += jobs[t].sum;
        if (sumsq) *sumsq += jobs[t].sumsq;
    }
    free(jobs);
    free(tid);
}

/* ---- PRBS --------------------------------------------------------------- */

/* prbs.py:14 */
uint32_t orc_prbs_tap(uint32_t k)
{
    switch (k) {
    case 7: return 6;  case 9: return 5;   case 11: return 9;  case 15: return 14;
    case 20: return 3; case 23: return 18; case 31: return 28;
    default: return 0;
    }
}

/* prbs.py:112-113 */
static inline uint32_t prbs_clock(uint32_t k, uint32_t tap, uint32_t *lfsr)
{
    uint32_t bit = ((*lfsr >> (k - 1)) ^ (*lfsr >> (tap - 1))) & 1;
    *lfsr = ((*lfsr << 1) | bit) & (uint32_t)((1ull << k) - 1);
    return bit;
}

void orc_prbs_bits(uint32_t k, uint32_t *state, uint64_t nbits, uint64_t *out)
{
    uint32_t tap = orc_prbs_tap(k);
    for (uint64_t i = 0; i < nbits; i++) {
        if ((i & 63) == 0) out[i >> 6] = 0;
        out[i >> 6] |= (uint64_t)prbs_clock(k, tap, state) << (i & 63);
    }
}

uint32_t orc_prbs_advance(uint32_t k, uint32_t state0, uint64_t steps)
{
    uint32_t tap = orc_prbs_tap(k), s = state0;
    for (uint64_t i = 0; i < steps; i++) prbs_clock(k, tap, &s);
    return s;
}

/* NEW definition: splitmix64 finaliser over (seed, word, draw). */
static inline uint64_t mix64(uint64_t z)
{
    z ^= z >> 30; z *= 0xBF58476D1CE4E5B9ull;
    z ^= z >> 27; z *= 0x94D049BB133111EBull;
    z ^= z >> 31;
    return z;
}

uint64_t orc_err_word(uint64_t seed, uint64_t W, uint32_t m)
{
    uint64_t e = ~(uint64_t)0;
    for (uint32_t j = 0; j < m; j++) e &= mix64(seed + ((W << 4) | j) * 0x9E3779B97F4A7C15ull);
    return e;
}

/* ---- PRBSErrorDetector, prbs.py:61-99 ------------------------------------ */

void orc_det_reset(uint32_t k, orc_det_state *s)
{
    s->bit_in = 0;                                   /* Signal() reset 0       */
    s->prbs = 1;                                     /* :62  reset=1           */
    s->err_sr = (uint32_t)((1ull << k) - 1);         /* :80  reset=2**k - 1    */
    s->reload_ctr = 0;                               /* :90                    */
}

uint32_t orc_det_clock(uint32_t k, orc_det_state *s, uint32_t bit)
{
    uint32_t tap = orc_prbs_tap(k);
    uint32_t mask = (uint32_t)((1ull << k) - 1);
    /* combinational values before the edge */
    uint32_t fb = ((s->prbs >> (k - 1)) ^ (s->prbs >> (tap - 1))) & 1;   /* :67 */
    uint32_t reload = s->reload_ctr != 0;                                 /* :99 */
    uint32_t prbs_in = reload ? s->bit_in : fb;                           /* :75-76 */
    uint32_t err = s->bit_in != fb;                                       /* :79 */
    uint32_t err_count = (uint32_t)__builtin_popcount(s->err_sr);         /* :84-86 */
    uint32_t fired = err_count > k / 2;
    /* clock edge: every right-hand side uses the pre-edge values */
    orc_det_state nx;
    nx.bit_in = bit & 1;                                                  /* :66 */
    nx.prbs = ((s->prbs << 1) | prbs_in) & mask;                          /* :68 */
    nx.err_sr = ((s->err_sr << 1) | err) & mask;                          /* :81 */
    nx.reload_ctr = s->reload_ctr;
    if (fired) {                                                          /* :91-98 */
        nx.reload_ctr = k + k / 2;
        nx.err_sr = 0;
    } else if (reload) {
        nx.reload_ctr = s->reload_ctr - 1;
    }
    *s = nx;
    /* outputs sampled after the edge, as the reference test does (:150-151) */
    uint32_t fb2 = ((s->prbs >> (k - 1)) ^ (s->prbs >> (tap - 1))) & 1;
    uint32_t err2 = s->bit_in != fb2;
    uint32_t reload2 = s->reload_ctr != 0;
    return err2 | (reload2 << 1) | (fired << 2);
}

void orc_prbs_detect(uint32_t k, uint32_t *gen_state, orc_det_state *det,
                     uint64_t first_bit, uint64_t nbits, uint32_t m, uint64_t seed,
                     orc_det_counters *ctr)
{
    uint32_t tap = orc_prbs_tap(k);
    uint64_t ew = 0;
    for (uint64_t i = 0; i < nbits; i++) {
        uint64_t abs_i = first_bit + i;
        if (m && (i == 0 || (abs_i & 63) == 0)) ew = orc_err_word(seed, abs_i >> 6, m);
        uint32_t e = m ? (uint32_t)((ew >> (abs_i & 63)) & 1) : 0;
        uint32_t bit = prbs_clock(k, tap, gen_state) ^ e;
        uint32_t o = orc_det_clock(k, det, bit);
        ctr->injected += e;
        ctr->err += o & 1;
        ctr->err_valid += (o & 1) & !((o >> 1) & 1);
        ctr->reload_cycles += (o >> 1) & 1;
        ctr->reload_events += (o >> 2) & 1;
    }
}

void orc_det_run_bits(uint32_t k, orc_det_state *det, const uint64_t *rx, const uint64_t *tx,
                      uint64_t nbits, orc_det_counters *ctr)
{
    for (uint64_t i = 0; i < nbits; i++) {
        uint32_t bit = (uint32_t)((rx[i >> 6] >> (i & 63)) & 1);
        uint32_t o = orc_det_clock(k, det, bit);
        if (tx) ctr->injected += (uint32_t)((tx[i >> 6] >> (i & 63)) & 1) ^ bit;
        ctr->err += o & 1;
        ctr->err_valid += (o & 1) & !((o >> 1) & 1);
        ctr->reload_cycles += (o >> 1) & 1;
        ctr->reload_events += (o >> 2) & 1;
    }
}
