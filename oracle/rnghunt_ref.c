/*
 * oracle/rnghunt_ref.c -- CPU ORACLE (test infrastructure, NOT the product path).
 *
 * The rnghunt worker loop, software/rnghunt/src/bin/rnghunt.rs:23-57, over the
 * restated GF(2) library (gf2.c).  The reference draws candidates from an
 * OS-seeded thread_rng (binary_matrix.rs:84; parity unpinned), so the build
 * DEFINES a reproducible candidate stream: candidate c of seed s is the matrix
 * binary_matrix.rs:81-101 produces when its draws come from the counter-based
 * generator below.  The GPU path (basebandboard_b200/csrc/rnghunt.cu) restates
 * the same definition independently.
 */
#include "rnghunt_ref.h"

#include <pthread.h>
#include <stdlib.h>
#include <string.h>

static inline uint64_t mix64(uint64_t z)
{
    z ^= z >> 30; z *= 0xBF58476D1CE4E5B9ull;
    z ^= z >> 27; z *= 0x94D049BB133111EBull;
    z ^= z >> 31;
    return z;
}

typedef struct { uint64_t base, i; } cand_rng;

/* NEW definition: draw #i of candidate (seed, c), uniform-ish index in [0, m). */
static uint32_t cand_draw(void *ctx, uint32_t m)
{
    cand_rng *g = (cand_rng *)ctx;
    g->i += 1;
    uint64_t z = mix64(g->base + g->i * 0x9E3779B97F4A7C15ull);
    return (uint32_t)(((z >> 32) * (uint64_t)m) >> 32);
}

static const size_t ORC_ROWWEIGHTS[8] = { 3, 4, 4, 4, 4, 4, 4, 4 }; /* rnghunt.rs:25 */

bmat orc_rnghunt_candidate(uint32_t n, uint64_t seed, uint64_t c)
{
    cand_rng g = { mix64(seed * 0x9E3779B97F4A7C15ull + c), 0 };
    return bm_random(n, n, ORC_ROWWEIGHTS, 8, cand_draw, &g);
}

/* rnghunt.rs:27-44 on one matrix.  Returns the stage the candidate reached. */
int orc_rnghunt_eval(const bmat *a, uint64_t *poly_words, uint32_t poly_cap_words)
{
    size_t n = a->nrows;
    uint8_t *ones = (uint8_t *)malloc(n);
    memset(ones, 1, n);
    bvec b = bv_from_bits(ones, n);                     /* :28 */
    bvec x = bm_recur(a, &b, 2 * n);                    /* :29 */
    uint8_t *xbits = (uint8_t *)malloc(2 * n);
    bv_to_bits(&x, xbits);                              /* :31 */
    for (size_t i = 0; i < n; i++) {                    /* :32 reverse */
        uint8_t t = xbits[i]; xbits[i] = xbits[2 * n - 1 - i]; xbits[2 * n - 1 - i] = t;
    }
    bvec xr = bv_from_bits(xbits, 2 * n);               /* :33 */
    bpoly p = gf2_berlekamp_massey(&xr);                /* :36 */
    int stage;
    if (poly_words) {
        memset(poly_words, 0, poly_cap_words * sizeof(uint64_t));
        size_t w = p.c.len < poly_cap_words ? p.c.len : poly_cap_words;
        memcpy(poly_words, p.c.data, w * sizeof(uint64_t));
    }
    if (bp_degree(&p) != (long)n) {                     /* :39-41 */
        stage = ORC_STAGE_DEGREE;
    } else {
        /* binary_polynomial.rs:178-212 unrolled so the stage is visible */
        if (!bv_get(&p.c, p.c.n - 1) || bv_count_ones(&p.c) % 2 != 1) {
            stage = ORC_STAGE_FILTER;
        } else {
            bvec f[64];
            size_t nf = gf2_get_factors(n, f, 64);
            if (!bp_check_integer(&p, &f[0])) {
                stage = ORC_STAGE_ORDER;
            } else {
                stage = ORC_STAGE_PRIMITIVE;
                for (size_t i = 1; i < nf; i++)
                    if (bp_check_integer(&p, &f[i])) { stage = ORC_STAGE_FACTOR; break; }
            }
            for (size_t i = 0; i < nf; i++) bv_free(&f[i]);
        }
        /* the unrolled verdict must be the library's verdict */
        if ((stage == ORC_STAGE_PRIMITIVE) != bp_is_primitive(&p)) abort();
    }
    bp_free(&p); bv_free(&xr); bv_free(&x); bv_free(&b);
    free(xbits); free(ones);
    return stage;
}

typedef struct {
    uint32_t n; uint64_t seed, first, count; int tid, nthreads;
    const uint64_t *list;          /* candidate indices, or NULL for first + i */
    uint8_t *stage_out; uint64_t stages[ORC_NSTAGES];
} sweep_job;

static void *sweep_worker(void *arg)
{
    sweep_job *j = (sweep_job *)arg;
    for (uint64_t i = (uint64_t)j->tid; i < j->count; i += (uint64_t)j->nthreads) {
        bmat a = orc_rnghunt_candidate(j->n, j->seed, j->list ? j->list[i] : j->first + i);
        int st = orc_rnghunt_eval(&a, NULL, 0);
        bm_free(&a);
        j->stages[st] += 1;
        if (j->stage_out) j->stage_out[i] = (uint8_t)st;
    }
    return NULL;
}

static void sweep_run(uint32_t n, uint64_t seed, uint64_t first, const uint64_t *list, uint64_t count,
                      uint8_t *stage_out, uint64_t *stage_counts, int threads);

void orc_rnghunt_sweep(uint32_t n, uint64_t seed, uint64_t first, uint64_t count,
                       uint8_t *stage_out, uint64_t *stage_counts, int threads)
{
    sweep_run(n, seed, first, NULL, count, stage_out, stage_counts, threads);
}

void orc_rnghunt_eval_list(uint32_t n, uint64_t seed, const uint64_t *cands, uint64_t count,
                           uint8_t *stage_out, uint64_t *stage_counts, int threads)
{
    sweep_run(n, seed, 0, cands, count, stage_out, stage_counts, threads);
}

static void sweep_run(uint32_t n, uint64_t seed, uint64_t first, const uint64_t *list, uint64_t count,
                      uint8_t *stage_out, uint64_t *stage_counts, int threads)
{
    if (threads < 1) threads = 1;
    sweep_job *jobs = (sweep_job *)calloc((size_t)threads, sizeof(sweep_job));
    pthread_t *tid = (pthread_t *)calloc((size_t)threads, sizeof(pthread_t));
    for (int t = 0; t < threads; t++) {
        sweep_job *j = &jobs[t];
        j->n = n; j->seed = seed; j->first = first; j->count = count;
        j->tid = t; j->nthreads = threads; j->stage_out = stage_out; j->list = list;
        pthread_create(&tid[t], NULL, sweep_worker, j);
    }
    for (int t = 0; t < threads; t++) {
        pthread_join(tid[t], NULL);
        if (stage_counts) for (int s = 0; s < ORC_NSTAGES; s++) stage_counts[s] += jobs[t].stages[s];
    }
    free(jobs);
    free(tid);
}
