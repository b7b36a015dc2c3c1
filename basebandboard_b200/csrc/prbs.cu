// prbs.cu -- PRBS generator and PRBSErrorDetector on the device (gateware/bbb/prbs.py).
//
//   PRBS (prbs.py:23-35, model :112-113): k-bit state s (reset 1), output bit
//   b = s[k-1] ^ s[tap-1], s' = ((s << 1) | b) & (2^k - 1).  The state at bit position p is the
//   window of the last k output bits, s_p[i] = b[p-1-i], and b[t] = b[t-k] ^ b[t-tap].
//
// Two facts carry the device design (DESIGN.md "PRBS path"):
//   * squaring the characteristic polynomial over GF(2) six times gives the same recurrence on
//     64-bit WORDS of the packed stream, W[j] = W[j-k] ^ W[j-tap] (and on any power-of-two word
//     stride), so the steady state is one 64-bit XOR per stored word: HBM-write bound;
//   * any position is reached by GF(2) jump-ahead: the host precomputes the k x k matrices M^(2^e)
//     of the one-bit step once per k, a thread applies those selected by the bits of its offset.
//
// The detector (prbs.py:43-99) is a sequential state machine with data-dependent reloads.  The bit
// range is cut into chunks, one thread each; a chunk first runs from the state the detector has
// when it is locked and no reload is pending (a function of the k+1 preceding channel bits only),
// then chunks whose predecessor ended in a different state are re-run from that state until
// nothing changes (fixed point == the sequential result, exactly).  Inside a chunk, 64-cycle
// words during which the detector provably stays locked and cannot fire are retired with a few
// word-wide operations; everything else is clocked bit by bit.
#include "bbb_internal.h"

// ---------------------------------------------------------------------------
// tables
// ---------------------------------------------------------------------------
struct PrbsTables {
    uint32_t pw[64][32];   // pw[e][i] = M^(2^e) . e_i : columns of the jump matrices
    uint64_t col64[32];    // the 64 output bits that follow state e_i (bit t of the word = b[p+t])
};

static inline uint32_t host_prbs_step(uint32_t k, uint32_t tap, uint32_t s)
{
    const uint32_t b = ((s >> (k - 1)) ^ (s >> (tap - 1))) & 1u;   // prbs.py:29 / :112
    return ((s << 1) | b) & (uint32_t)((1ull << k) - 1);           // prbs.py:30 / :113
}

// the state one clock earlier (the LFSR step is a bijection)
static inline uint32_t host_prbs_unstep(uint32_t k, uint32_t tap, uint32_t s)
{
    const uint32_t hi = (s ^ (s >> tap)) & 1u;   // s_prev[k-1] = b ^ s_prev[tap-1] = s[0] ^ s[tap]
    return (s >> 1) | (hi << (k - 1));
}

static inline uint32_t host_matvec(const uint32_t *cols, uint32_t k, uint32_t x)
{
    uint32_t y = 0;
    for (uint32_t i = 0; i < k; i++)
        if ((x >> i) & 1u) y ^= cols[i];
    return y;
}

static void build_tables(uint32_t k, uint32_t tap, PrbsTables &t)
{
    memset(&t, 0, sizeof t);
    for (uint32_t i = 0; i < k; i++) t.pw[0][i] = host_prbs_step(k, tap, 1u << i);
    for (uint32_t e = 1; e < 64; e++)
        for (uint32_t i = 0; i < k; i++) t.pw[e][i] = host_matvec(t.pw[e - 1], k, t.pw[e - 1][i]);
    for (uint32_t i = 0; i < k; i++) {
        uint32_t s = 1u << i;
        uint64_t w = 0;
        for (uint32_t b = 0; b < 64; b++) {
            w |= (uint64_t)(((s >> (k - 1)) ^ (s >> (tap - 1))) & 1u) << b;
            s = host_prbs_step(k, tap, s);
        }
        t.col64[i] = w;
    }
}

static uint32_t host_jump(const PrbsTables &t, uint32_t k, uint32_t s, uint64_t delta)
{
    for (uint32_t e = 0; delta; e++, delta >>= 1)
        if (delta & 1) s = host_matvec(t.pw[e], k, s);
    return s;
}

static int prbs_slot(uint32_t k)
{
    switch (k) {
    case 7: return 0; case 9: return 1; case 11: return 2; case 15: return 3;
    case 20: return 4; case 23: return 5; case 31: return 6; default: return -1;
    }
}

// Host + device copies of the tables for k, built on first use and kept by the ctx.
static int get_tables(bbb_ctx *c, uint32_t k, const PrbsTables **host, const PrbsTables **dev)
{
    const int slot = prbs_slot(k);
    if (slot < 0) return c->fail(BBB_EINVAL, "k=%u is not one of prbs.py:14's TAPS keys", k);
    if (!c->prbs_host[slot]) {
        PrbsTables *h = new (std::nothrow) PrbsTables;
        if (!h) return c->fail(BBB_ENOMEM, "out of memory");
        build_tables(k, bbb_prbs_tap(k), *h);
        void *d = nullptr;
        cudaError_t e = cudaMalloc(&d, sizeof(PrbsTables));
        if (e == cudaSuccess) e = cudaMemcpy(d, h, sizeof(PrbsTables), cudaMemcpyHostToDevice);
        if (e != cudaSuccess) {
            if (d) cudaFree(d);
            delete h;
            return c->cuda_fail(e, "upload PRBS tables");
        }
        c->prbs_host[slot] = h;
        c->prbs_dev[slot] = d;
    }
    *host = (const PrbsTables *)c->prbs_host[slot];
    *dev = (const PrbsTables *)c->prbs_dev[slot];
    return BBB_OK;
}

void bbb_prbs_free_tables(bbb_ctx *c)
{
    for (int i = 0; i < 8; i++) {
        if (c->prbs_host[i]) delete (PrbsTables *)c->prbs_host[i];
        if (c->prbs_dev[i]) cudaFree(c->prbs_dev[i]);
        c->prbs_host[i] = nullptr;
        c->prbs_dev[i] = nullptr;
    }
}

// ---------------------------------------------------------------------------
// device helpers
// ---------------------------------------------------------------------------
struct PrbsShared {
    uint32_t pw[64][32];
    uint64_t col64[32];
};

template <int K>
__device__ __forceinline__ void load_tables(PrbsShared &sh, const PrbsTables *__restrict__ t)
{
    const uint32_t *src = (const uint32_t *)t;
    uint32_t *dst = (uint32_t *)&sh;
    for (uint32_t i = threadIdx.x; i < sizeof(PrbsTables) / 4; i += blockDim.x) dst[i] = src[i];
    __syncthreads();
}

template <int K>
__device__ __forceinline__ uint32_t matvec(const uint32_t *cols, uint32_t x)
{
    uint32_t y = 0;
#pragma unroll
    for (int i = 0; i < K; i++) y ^= (0u - ((x >> i) & 1u)) & cols[i];
    return y;
}

template <int K>
__device__ __forceinline__ uint32_t jump(const PrbsShared &sh, uint32_t s, uint64_t delta)
{
    for (uint32_t e = 0; delta; e++, delta >>= 1)
        if (delta & 1) s = matvec<K>(sh.pw[e], s);
    return s;
}

// the 64 bits b[p .. p+63] that follow state s_p
template <int K>
__device__ __forceinline__ uint64_t word_of(const PrbsShared &sh, uint32_t s)
{
    uint64_t w = 0;
#pragma unroll
    for (int i = 0; i < K; i++) w ^= (0ull - (uint64_t)((s >> i) & 1u)) & sh.col64[i];
    return w;
}

// s_{p+64} from the word b[p .. p+63]: s[i] = b[p+63-i]
template <int K>
__device__ __forceinline__ uint32_t state_after(uint64_t w)
{
    return __brev((uint32_t)(w >> 32)) & (uint32_t)((1ull << K) - 1);
}

// ---------------------------------------------------------------------------
// stream kernel: packed bits, one warp per chunk, lane l owns words VEC*l .. VEC*l+VEC-1 of
// every group of 32*VEC words; K groups of history stay in registers.
// ---------------------------------------------------------------------------
template <int K, int TAP, int VEC>
__global__ void __launch_bounds__(128)
prbs_stream_kernel(const PrbsTables *__restrict__ tab, uint32_t s_first, uint64_t nwords, uint32_t reps,
                   uint64_t *__restrict__ out)
{
    __shared__ PrbsShared sh;
    load_tables<K>(sh, tab);
    constexpr uint32_t GW = 32 * VEC;                    // words per warp step
    constexpr uint32_t LOG_GW = VEC == 2 ? 6 : 5;
    const uint32_t lane = threadIdx.x & 31;
    const uint64_t warp = ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const uint64_t chunk_words = (uint64_t)GW * K * reps;
    const uint64_t cbase = warp * chunk_words;
    if (cbase >= nwords) return;
    uint64_t h[K][VEC];
    {
        // History = the chunk's first K*GW words.  Only the first K come from jump-ahead (lane l: word l); the
        // rest follow from the word recurrence W[j] = W[j - K*2^e] ^ W[j - TAP*2^e], valid for every power of two
        // (Frobenius), applied in the warp's shared buffer with the largest stride the known prefix allows, so a
        // pass extends the prefix by TAP*2^e words at once (a handful of passes instead of K matrix products
        // and 2K word look-ups per lane: the seeding was 0.10 ms of a 64 MB call).
        extern __shared__ uint64_t prbs_wbuf[];
        uint64_t *const wb = prbs_wbuf + (size_t)(threadIdx.x >> 5) * (K * GW);
        if (lane < K) wb[lane] = word_of<K>(sh, jump<K>(sh, s_first, (cbase + lane) << 6));
        __syncwarp();
        constexpr uint32_t TOTAL = K * GW;
        uint32_t have = K, e = 0;
        while (have < TOTAL) {
            while (((uint32_t)K << (e + 1)) <= have) e++;
            const uint32_t dk = (uint32_t)K << e, dt = (uint32_t)TAP << e;
            const uint32_t batch = dt < TOTAL - have ? dt : TOTAL - have;
            for (uint32_t i = lane; i < batch; i += 32) wb[have + i] = wb[have + i - dk] ^ wb[have + i - dt];
            __syncwarp();
            have += batch;
        }
#pragma unroll
        for (int m = 0; m < K; m++) {
#pragma unroll
            for (int v = 0; v < VEC; v++) h[m][v] = wb[m * GW + VEC * lane + v];
        }
    }
    uint64_t j = cbase + (uint64_t)VEC * lane;
    const bool whole = cbase + chunk_words <= nwords;     // warp-uniform
    for (uint32_t r = 0; r < reps; r++) {
#pragma unroll
        for (int m = 0; m < K; m++) {
            if (r) {
#pragma unroll
                for (int v = 0; v < VEC; v++) h[m][v] ^= h[(m + K - TAP) % K][v];   // W[j] = W[j-k] ^ W[j-tap]
            }
            if (VEC == 2) {
                if (whole || j + 1 < nwords)
                    __stcs((ulonglong2 *)(out + j), make_ulonglong2(h[m][0], h[m][VEC - 1]));
                else if (j < nwords)
                    out[j] = h[m][0];
            } else {
                if (whole || j < nwords) __stcs(out + j, h[m][0]);
            }
            j += GW;
        }
    }
}

// ---------------------------------------------------------------------------
// detector
// ---------------------------------------------------------------------------
struct DetectParams {
    const PrbsTables *tab;
    uint32_t s_m1;             // generator state at bit first_bit - 1
    uint32_t err_m;
    uint64_t seed, first_bit, nbits;
    uint32_t chunk_words, nchunks, round;
    uint4 det0;                // detector state on entry (chunk 0)
    uint4 *start_used;         // [nchunks]
    uint4 *end_prev, *end_cur; // [nchunks] double-buffered end states
    unsigned long long *ctr;   // [5][nchunks]
    uint32_t *changed;
    const uint64_t *err_words; // caller-supplied channel errors (word j = bits [64j, 64j+64) after first_bit), or null
};

// DESIGN.md "channel": Bernoulli(2^-m) error word for absolute bits [64W, 64W+64): the AND of m
// splitmix64-finalised counters keyed (seed, W, draw).
__device__ __forceinline__ uint64_t mix64(uint64_t z)
{
    z ^= z >> 30; z *= 0xBF58476D1CE4E5B9ull;
    z ^= z >> 27; z *= 0x94D049BB133111EBull;
    z ^= z >> 31;
    return z;
}

__device__ __forceinline__ uint64_t err_word(uint64_t seed, uint64_t W, uint32_t m)
{
    uint64_t e = ~0ull;
    for (uint32_t j = 0; j < m; j++) e &= mix64(seed + ((W << 4) | j) * 0x9E3779B97F4A7C15ull);
    return e;
}

// error bits of the 64 channel bits starting at absolute bit `a`
__device__ __forceinline__ uint64_t err_bits_at(uint64_t seed, uint64_t a, uint32_t m)
{
    if (!m) return 0;
    const uint32_t sh = (uint32_t)(a & 63);
    uint64_t e = err_word(seed, a >> 6, m) >> sh;
    if (sh) e |= err_word(seed, (a >> 6) + 1, m) << (64 - sh);
    return e;
}

template <int K, int TAP>
__global__ void __launch_bounds__(128)
prbs_detect_kernel(DetectParams p)
{
    __shared__ PrbsShared sh;
    __shared__ uint64_t ring[K][128];
    load_tables<K>(sh, p.tab);
    const uint32_t c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= p.nchunks) return;
    constexpr uint32_t MASK = (uint32_t)((1ull << K) - 1);
    const uint32_t tid = threadIdx.x;

    const uint64_t w0 = (uint64_t)c * p.chunk_words;              // first word of the chunk
    const uint64_t bit0 = w0 << 6;                                 // relative to first_bit
    uint64_t bits_left = p.nbits - bit0;
    if (bits_left > ((uint64_t)p.chunk_words << 6)) bits_left = (uint64_t)p.chunk_words << 6;

    // ---- which start state? ----
    uint4 start;
    bool have_start = false;
    if (c == 0) { start = p.det0; have_start = true; }
    else if (p.round > 0) { start = p.end_prev[c - 1]; have_start = true; }
    if (p.round > 0) {
        const uint4 u = p.start_used[c];
        if (u.x == start.x && u.y == start.y && u.z == start.z && u.w == start.w) {
            p.end_cur[c] = p.end_prev[c];
            return;
        }
        *p.changed = 1u;
    }

    // ---- generator: states either side of the chunk start, first K words ----
    const uint32_t s_pm1 = jump<K>(sh, p.s_m1, bit0);
    const uint32_t s_p = matvec<K>(sh.pw[0], s_pm1);
    {
        uint32_t s = s_p;
#pragma unroll 1
        for (int m = 0; m < K; m++) {
            const uint64_t w = word_of<K>(sh, s);
            ring[m][tid] = w;
            s = state_after<K>(w);
        }
    }
    // "previous word" whose top K+1 bits encode b[p-1] (bit 63) and s_{p-1}[i] (bit 62-i)
    uint64_t wprev = __brevll(((uint64_t)s_pm1 << 1) | (uint64_t)(s_p & 1u));

    if (!have_start) {
        // locked, nothing pending: bit_in = received b[p-1], prbs = s_{p-1}, err_sr = the error bits
        // of the K cycles before, reload_ctr = 0
        const uint64_t eprev = p.err_words ? p.err_words[w0 - 1] : err_bits_at(p.seed, p.first_bit + bit0 - 64, p.err_m);
        start.x = (uint32_t)((wprev ^ eprev) >> 63);
        start.y = s_pm1;
        start.z = (uint32_t)__brevll(eprev << 1) & MASK;
        start.w = 0;
    }
    p.start_used[c] = start;
    uint32_t bit_in = start.x, prbs = start.y, err_sr = start.z, ctr = start.w;
    unsigned long long n_err = 0, n_valid = 0, n_fire = 0, n_rcyc = 0, n_inj = 0;

    uint32_t pos = 0, tpos = K - TAP;    // ring slots of W[j-K] and W[j-TAP]
    const uint64_t nwords = (bits_left + 63) >> 6;
    for (uint64_t j = 0; j < nwords; j++) {
        uint64_t w;
        if (j < (uint64_t)K) w = ring[pos][tid];
        else { w = ring[pos][tid] ^ ring[tpos][tid]; ring[pos][tid] = w; }
        pos = pos + 1 == K ? 0 : pos + 1;
        tpos = tpos + 1 == K ? 0 : tpos + 1;

        const uint32_t nb = bits_left - (j << 6) >= 64 ? 64u : (uint32_t)(bits_left - (j << 6));
        uint64_t e = p.err_words ? p.err_words[w0 + j] : err_bits_at(p.seed, p.first_bit + bit0 + (j << 6), p.err_m);
        if (nb < 64) e &= (1ull << nb) - 1;
        n_inj += __popcll(e);
        const uint64_t r = w ^ e;                                     // received bits

        const uint32_t prbs_exp = __brev((uint32_t)(wprev >> 31)) & MASK;
        const uint32_t pend = bit_in ^ (uint32_t)(wprev >> 63);
        const uint32_t worst = __popc(err_sr) + pend + __popcll(e & 0x7fffffffffffffffull);
        if (nb == 64 && ctr == 0 && prbs == prbs_exp && worst <= K / 2) {
            // locked for the whole word and no window of K cycles can exceed K/2 errors: err pulses
            // are exactly the channel errors, no reload
            const uint32_t ne = __popcll(e);
            n_err += ne;
            n_valid += ne;
            err_sr = (uint32_t)__brevll(e << 1) & MASK;
            bit_in = (uint32_t)(r >> 63);
            prbs = __brev((uint32_t)(w >> 31)) & MASK;
        } else {
            for (uint32_t i = 0; i < nb; i++) {
                // combinational values before the edge (prbs.py:67,75-76,79,84-86,99)
                const uint32_t fb = ((prbs >> (K - 1)) ^ (prbs >> (TAP - 1))) & 1u;
                const uint32_t reload = ctr != 0;
                const uint32_t prbs_in = reload ? bit_in : fb;
                const uint32_t err = bit_in ^ fb;
                const bool fired = __popc(err_sr) > K / 2;
                // the edge (prbs.py:66,68,81,91-98)
                prbs = ((prbs << 1) | prbs_in) & MASK;
                err_sr = ((err_sr << 1) | err) & MASK;
                bit_in = (uint32_t)(r >> i) & 1u;
                if (fired) { ctr = K + K / 2; err_sr = 0; n_fire++; }
                else if (reload) ctr--;
                // outputs after the edge, as prbs.py:150-151 samples them
                const uint32_t fb2 = ((prbs >> (K - 1)) ^ (prbs >> (TAP - 1))) & 1u;
                const uint32_t err2 = bit_in ^ fb2;
                const uint32_t rl2 = ctr != 0;
                n_err += err2;
                n_valid += err2 & (rl2 ^ 1u);
                n_rcyc += rl2;
            }
        }
        wprev = w;
    }
    p.end_cur[c] = make_uint4(bit_in, prbs, err_sr, ctr);
    p.ctr[0 * (size_t)p.nchunks + c] = n_err;
    p.ctr[1 * (size_t)p.nchunks + c] = n_valid;
    p.ctr[2 * (size_t)p.nchunks + c] = n_fire;
    p.ctr[3 * (size_t)p.nchunks + c] = n_rcyc;
    p.ctr[4 * (size_t)p.nchunks + c] = n_inj;
}

// out[0..4] = column sums of ctr[5][n]; out[5..8] = end state of the last chunk
__global__ void __launch_bounds__(256)
prbs_detect_reduce_kernel(const unsigned long long *__restrict__ ctr, const uint4 *__restrict__ end, uint32_t n,
                          unsigned long long *__restrict__ out)
{
    __shared__ unsigned long long s[256];
    const uint32_t q = blockIdx.x;   // counter index
    unsigned long long a = 0;
    for (uint32_t i = threadIdx.x; i < n; i += blockDim.x) a += ctr[(size_t)q * n + i];
    s[threadIdx.x] = a;
    __syncthreads();
    for (uint32_t st = 128; st; st >>= 1) {
        if (threadIdx.x < st) s[threadIdx.x] += s[threadIdx.x + st];
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        out[q] = s[0];
        if (q == 0) {
            const uint4 e = end[n - 1];
            out[5] = e.x; out[6] = e.y; out[7] = e.z; out[8] = e.w;
        }
    }
}

// ---------------------------------------------------------------------------
// C ABI
// ---------------------------------------------------------------------------
template <int VEC>
static int launch_stream(bbb_ctx *c, uint32_t k, const PrbsTables *dtab, uint32_t s_first, uint64_t nwords,
                         uint64_t *d_out)
{
    // one warp per chunk of 32*VEC*k*reps words; a single wave of sm_count*12 warps when the range is
    // long enough, so every warp amortises its jump-ahead over the same number of stores
    const uint64_t unit = (uint64_t)32 * VEC * k;
    uint64_t reps = (nwords + unit * c->sm_count * 12 - 1) / (unit * c->sm_count * 12);
    if (c->tune_prbs_stream_reps) reps = c->tune_prbs_stream_reps;
    if (reps < 1) reps = 1;
    if (reps > 1u << 20) reps = 1u << 20;
    const uint64_t warps = (nwords + unit * reps - 1) / (unit * reps);
    const uint64_t blocks = (warps + 3) / 4;
    if (blocks > 0x7fffffffull) return c->fail(BBB_EINVAL, "range too large");
#define PRBS_CASE(K, TAP) \
    case K: \
        BBB_CUDA(c, cudaFuncSetAttribute(prbs_stream_kernel<K, TAP, VEC>, cudaFuncAttributeMaxDynamicSharedMemorySize, \
                                         4 * K * 32 * VEC * 8)); \
        prbs_stream_kernel<K, TAP, VEC><<<(uint32_t)blocks, 128, 4 * K * 32 * VEC * 8, c->stream>>>(dtab, s_first, nwords, \
                                                                                                     (uint32_t)reps, d_out); break
    switch (k) {
        PRBS_CASE(7, 6); PRBS_CASE(9, 5); PRBS_CASE(11, 9); PRBS_CASE(15, 14);
        PRBS_CASE(20, 3); PRBS_CASE(23, 18); PRBS_CASE(31, 28);
    default: return c->fail(BBB_EINVAL, "bad k");
    }
#undef PRBS_CASE
    c->last_launches++;
    BBB_CUDA(c, cudaGetLastError());
    return BBB_OK;
}

static int check_prbs_args(bbb_ctx *c, uint32_t k, uint32_t state0)
{
    if (!c) return BBB_EINVAL;
    if (!bbb_prbs_tap(k)) return c->fail(BBB_EINVAL, "k=%u is not one of prbs.py:14's TAPS keys", k);
    if (state0 >> k) return c->fail(BBB_EINVAL, "state0 has more than k bits");
    return BBB_OK;
}

extern "C" {

uint32_t bbb_prbs_tap(uint32_t k)
{
    switch (k) {   // prbs.py:14
    case 7: return 6; case 9: return 5; case 11: return 9; case 15: return 14;
    case 20: return 3; case 23: return 18; case 31: return 28; default: return 0;
    }
}

int bbb_prbs_stream_dev(bbb_ctx *c, uint32_t k, uint32_t state0, uint64_t first_bit, uint64_t nbits,
                        uint64_t *out_words_dev)
{
    int rc = check_prbs_args(c, k, state0);
    if (rc) return rc;
    if (nbits % 64) return c->fail(BBB_EINVAL, "nbits must be a multiple of 64");
    if (!out_words_dev && nbits) return c->fail(BBB_EINVAL, "out is null");
    BBB_CUDA(c, cudaSetDevice(c->device));
    const PrbsTables *ht, *dt;
    rc = get_tables(c, k, &ht, &dt);
    if (rc) return rc;
    c->begin_call();
    c->last_kernel = "prbs_stream";
    if (nbits) {
        const uint32_t s_first = host_jump(*ht, k, state0, first_bit);
        if (((uintptr_t)out_words_dev & 15) == 0) rc = launch_stream<2>(c, k, dt, s_first, nbits / 64, out_words_dev);
        else rc = launch_stream<1>(c, k, dt, s_first, nbits / 64, out_words_dev);
        if (rc) return rc;
    }
    c->end_call();
    return BBB_OK;
}

int bbb_prbs_stream(bbb_ctx *c, uint32_t k, uint32_t state0, uint64_t first_bit, uint64_t nbits,
                    uint64_t *out_words)
{
    int rc = check_prbs_args(c, k, state0);
    if (rc) return rc;
    if (nbits % 64) return c->fail(BBB_EINVAL, "nbits must be a multiple of 64");
    if (!out_words && nbits) return c->fail(BBB_EINVAL, "out is null");
    return bbb_stream_to_host(c, 4, out_words, (size_t)(nbits / 8), 8 * 4096, (size_t)64 << 20,
                              [&](size_t off, size_t bytes, void *dev) {
                                  return bbb_prbs_stream_dev(c, k, state0, first_bit + (uint64_t)off * 8,
                                                             (uint64_t)bytes * 8, (uint64_t *)dev);
                              });
}

// splitmix64-keyed channel, host copy (bbb_prbs_spec_state only)
static uint64_t host_mix64(uint64_t z)
{
    z ^= z >> 30; z *= 0xBF58476D1CE4E5B9ull;
    z ^= z >> 27; z *= 0x94D049BB133111EBull;
    z ^= z >> 31;
    return z;
}

static uint64_t host_err_bits_at(uint64_t seed, uint64_t a, uint32_t m)
{
    if (!m) return 0;
    uint64_t w[2];
    for (int q = 0; q < 2; q++) {
        uint64_t e = ~0ull;
        for (uint32_t j = 0; j < m; j++) e &= host_mix64(seed + ((((a >> 6) + q) << 4) | j) * 0x9E3779B97F4A7C15ull);
        w[q] = e;
    }
    const uint32_t sh = (uint32_t)(a & 63);
    return sh ? (w[0] >> sh) | (w[1] << (64 - sh)) : w[0];
}

static uint64_t host_brev64(uint64_t x)
{
    uint64_t y = 0;
    for (int i = 0; i < 64; i++) y |= ((x >> i) & 1ull) << (63 - i);
    return y;
}

int bbb_prbs_spec_state(bbb_ctx *c, uint32_t k, uint64_t first_bit, uint32_t err_m, uint64_t seed,
                        bbb_det_state *det)
{
    // pure host arithmetic: works without a ctx (and without a GPU) so that shard planning can run anywhere
    const uint32_t tap = bbb_prbs_tap(k);
    if (!tap || !det || err_m > 16 || first_bit < 64) {
        const char *why = !tap ? "k is not one of prbs.py:14's TAPS keys" : !det ? "det is null"
                          : err_m > 16 ? "err_m must be 0..16"
                          : "first_bit must be >= 64 (the detector starts from reset at 0)";
        return c ? c->fail(BBB_EINVAL, "%s", why) : BBB_EINVAL;
    }
    PrbsTables *t = new (std::nothrow) PrbsTables;
    if (!t) return BBB_ENOMEM;
    build_tables(k, tap, *t);
    const uint32_t mask = (uint32_t)((1ull << k) - 1);
    const uint32_t s_p = host_jump(*t, k, 1u, first_bit);
    delete t;
    const uint32_t s_pm1 = host_prbs_unstep(k, tap, s_p);
    const uint64_t eprev = host_err_bits_at(seed, first_bit - 64, err_m);
    det->bit_in = (s_p & 1u) ^ (uint32_t)(eprev >> 63);
    det->prbs = s_pm1;
    det->err_sr = (uint32_t)host_brev64(eprev << 1) & mask;
    det->reload_ctr = 0;
    return BBB_OK;
}

}  // extern "C"

// the detector over a bit range; the channel is the keyed Bernoulli source (err_words null) or the
// caller's error words on the device
static int detect_impl(bbb_ctx *c, uint32_t k, uint64_t first_bit, uint64_t nbits, uint32_t err_m, uint64_t seed,
                       const uint64_t *err_words, bbb_det_state *det, bbb_det_counters *ctr)
{
    int rc = check_prbs_args(c, k, 0);
    if (rc) return rc;
    if (!ctr) return c->fail(BBB_EINVAL, "ctr is null");
    if (err_m > 16) return c->fail(BBB_EINVAL, "err_m must be 0..16");
    if (nbits >> 62) return c->fail(BBB_EINVAL, "nbits too large");
    BBB_CUDA(c, cudaSetDevice(c->device));
    const PrbsTables *ht, *dt;
    rc = get_tables(c, k, &ht, &dt);
    if (rc) return rc;
    c->begin_call();
    c->last_kernel = "prbs_detect";
    c->last_rounds = 0;
    if (!nbits) { c->end_call(); return BBB_OK; }
    const uint32_t tap = bbb_prbs_tap(k), mask = (uint32_t)((1ull << k) - 1);
    DetectParams p;
    p.tab = dt;
    p.s_m1 = host_prbs_unstep(k, tap, host_jump(*ht, k, 1u, first_bit));
    p.err_m = err_m;
    p.seed = seed;
    p.err_words = err_words;
    p.first_bit = first_bit;
    p.nbits = nbits;
    const uint64_t nwords = (nbits + 63) / 64;
    uint64_t cw = (nwords + (uint64_t)c->sm_count * 1024 - 1) / ((uint64_t)c->sm_count * 1024);
    if (cw < 32) cw = 32;
    if (cw > 2048) cw = 2048;
    if (c->tune_prbs_chunk_words) cw = c->tune_prbs_chunk_words;
    uint64_t nchunks = (nwords + cw - 1) / cw;
    while (nchunks > 0x7fffff00ull) { cw *= 2; nchunks = (nwords + cw - 1) / cw; }
    p.chunk_words = (uint32_t)cw;
    p.nchunks = (uint32_t)nchunks;
    if (det) p.det0 = make_uint4(det->bit_in & 1u, det->prbs & mask, det->err_sr & mask, det->reload_ctr);
    else p.det0 = make_uint4(0u, 1u, mask, 0u);   // reset values, prbs.py:62,80,90
    // scratch: start_used | end[2] | ctr[5] | out[9] | changed
    const size_t n = p.nchunks;
    const size_t bytes = n * 16 * 3 + n * 8 * 5 + 9 * 8 + 16;
    char *d = (char *)c->scratch(5, bytes);
    if (!d) return BBB_ECUDA;
    p.start_used = (uint4 *)d;
    uint4 *end0 = (uint4 *)(d + n * 16), *end1 = (uint4 *)(d + n * 32);
    p.ctr = (unsigned long long *)(d + n * 48);
    unsigned long long *d_out = p.ctr + n * 5;
    p.changed = (uint32_t *)(d_out + 9);
    const uint32_t blocks = (p.nchunks + 127) / 128;
    uint32_t round = 0;
    for (;;) {
        p.round = round;
        p.end_prev = (round & 1) ? end0 : end1;
        p.end_cur = (round & 1) ? end1 : end0;
        BBB_CUDA(c, cudaMemsetAsync(p.changed, 0, 4, c->stream));
#define PRBS_CASE(K, TAP) case K: prbs_detect_kernel<K, TAP><<<blocks, 128, 0, c->stream>>>(p); break
        switch (k) {
            PRBS_CASE(7, 6); PRBS_CASE(9, 5); PRBS_CASE(11, 9); PRBS_CASE(15, 14);
            PRBS_CASE(20, 3); PRBS_CASE(23, 18); PRBS_CASE(31, 28);
        default: return c->fail(BBB_EINVAL, "bad k");
        }
#undef PRBS_CASE
        c->last_launches++;
        BBB_CUDA(c, cudaGetLastError());
        round++;
        if (round == 1) {
            if (p.nchunks == 1) break;
            continue;   // round 1 checks every chunk's guess against its predecessor
        }
        uint32_t changed = 0;
        BBB_CUDA(c, c->d2h(&changed, p.changed, 4));
        BBB_CUDA(c, cudaStreamSynchronize(c->stream));
        if (!changed) break;
        if (round > p.nchunks + 1) return c->fail(BBB_ESTATE, "detector fix-up did not converge");
    }
    c->last_rounds = round;
    prbs_detect_reduce_kernel<<<5, 256, 0, c->stream>>>(p.ctr, p.end_cur, p.nchunks, d_out);
    c->last_launches++;
    BBB_CUDA(c, cudaGetLastError());
    c->end_call();
    unsigned long long out[9];
    BBB_CUDA(c, c->d2h(out, d_out, sizeof out));
    BBB_CUDA(c, cudaStreamSynchronize(c->stream));
    ctr->err += out[0];
    ctr->err_valid += out[1];
    ctr->reload_events += out[2];
    ctr->reload_cycles += out[3];
    ctr->injected += out[4];
    if (det) {
        det->bit_in = (uint32_t)out[5];
        det->prbs = (uint32_t)out[6];
        det->err_sr = (uint32_t)out[7];
        det->reload_ctr = (uint32_t)out[8];
    }
    return BBB_OK;
}

// AWGN slicer channel (tx.py:75-81, rx.py:28): the transmitted sample of bit b is +A or -A plus
// noise = g * noise_var, registered as a 12-bit signed word (wraps like the gateware's Signal((12, True)));
// the receiver slices it at 0 (sliced = ~sample[-1]) and takes sample `delay` of the spb samples of a bit.
// One thread per 64 bits: err word = sliced bits XOR transmitted bits.
__global__ void __launch_bounds__(256)
slicer_err_kernel(const int8_t *__restrict__ g, const uint64_t *__restrict__ tx, uint64_t nwords, uint32_t spb,
                  uint32_t delay, int32_t amp, int32_t noise_var, uint64_t *__restrict__ err)
{
    const uint64_t j = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= nwords) return;
    const uint64_t w = tx[j];
    const int8_t *gp = g + (j * 64) * spb + delay;
    uint64_t rx = 0;
    if (spb == 1 && ((uintptr_t)g & 15) == 0) {          // decimated noise: 64 consecutive bytes, four 128-bit loads
#pragma unroll
        for (uint32_t q = 0; q < 4; q++) {
            const uint4 v = __ldcs((const uint4 *)gp + q);
            const uint32_t vv[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
            for (uint32_t i = 0; i < 16; i++) {
                const int32_t noise = (int32_t)(int8_t)((vv[i >> 2] >> (8 * (i & 3))) & 0xffu) * noise_var;
                const int32_t x = (((w >> (16 * q + i)) & 1ull) ? amp : -amp) + noise;
                const int32_t x12 = ((x + 2048) & 4095) - 2048;
                rx |= (uint64_t)(x12 >= 0) << (16 * q + i);
            }
        }
        err[j] = rx ^ w;
        return;
    }
#pragma unroll 4
    for (uint32_t i = 0; i < 64; i++) {
        const int32_t noise = (int32_t)gp[(size_t)i * spb] * noise_var;
        const int32_t x = (((w >> i) & 1ull) ? amp : -amp) + noise;
        const int32_t x12 = ((x + 2048) & 4095) - 2048;
        rx |= (uint64_t)(x12 >= 0) << i;
    }
    err[j] = rx ^ w;
}

extern "C" {

int bbb_prbs_detect(bbb_ctx *c, uint32_t k, uint64_t first_bit, uint64_t nbits, uint32_t err_m, uint64_t seed,
                    bbb_det_state *det, bbb_det_counters *ctr)
{
    return detect_impl(c, k, first_bit, nbits, err_m, seed, nullptr, det, ctr);
}

int bbb_prbs_detect_errs_dev(bbb_ctx *c, uint32_t k, uint64_t first_bit, uint64_t nbits, const uint64_t *err_words_dev,
                             bbb_det_state *det, bbb_det_counters *ctr)
{
    if (!c) return BBB_EINVAL;
    if (!err_words_dev) return c->fail(BBB_EINVAL, "err_words_dev is null");
    return detect_impl(c, k, first_bit, nbits, 0, 0, err_words_dev, det, ctr);
}

int bbb_prbs_slicer_detect(bbb_ctx *c, uint32_t k, uint64_t first_bit, uint64_t nbits, const uint64_t *x0,
                           uint32_t amplitude, uint32_t noise_var, uint32_t spb, uint32_t delay, bbb_det_state *det,
                           bbb_det_counters *ctr)
{
    int rc = check_prbs_args(c, k, 0);
    if (rc) return rc;
    if (!ctr || !x0) return c->fail(BBB_EINVAL, "null argument");
    if (nbits & 63) return c->fail(BBB_EINVAL, "nbits must be a multiple of 64");
    if (amplitude > 2047 || noise_var > 15) return c->fail(BBB_EINVAL, "amplitude 0..2047, noise_var 0..15 (tx.py:46-49)");
    if (!spb || (spb & (spb - 1)) || spb > 16 || delay >= spb)
        return c->fail(BBB_EINVAL, "spb must be a power of two <= 16 and delay < spb (rx.py:18-19)");
    if (c->n == 0 || c->n > 256) return c->fail(BBB_ESTATE, "load a recurrence with n <= 256 first");
    if ((first_bit + nbits) >> 58) return c->fail(BBB_EINVAL, "bit range too large");
    bbb_det_state d0;
    if (det) d0 = *det;
    else { d0.bit_in = 0; d0.prbs = 1; d0.err_sr = (uint32_t)((1ull << k) - 1); d0.reload_ctr = 0; }
    c->begin_call();
    struct Nest { bbb_ctx *c; Nest(bbb_ctx *cc) : c(cc) { c->nest++; } ~Nest() { c->nest--; } } nest(c);
    const uint64_t CH = c->tune_slicer_chunk_bits ? c->tune_slicer_chunk_bits : (1ull << 32);   // 4 GiB of sampled int8 noise per pass
    uint32_t rounds = 0;
    for (uint64_t off = 0; off < nbits; off += CH) {
        const uint64_t nb = nbits - off < CH ? nbits - off : CH, nw = nb / 64;
        int8_t *d_g = (int8_t *)c->scratch(9, (size_t)nb);
        uint64_t *d_tx = (uint64_t *)c->scratch(10, (size_t)nw * 16);
        if (!d_g || !d_tx) return BBB_ECUDA;
        uint64_t *d_err = d_tx + nw;
        rc = bbb_prbs_stream_dev(c, k, 1u, first_bit + off, nb, d_tx);
        if (rc) return rc;
        // only the sample the receiver keeps of each bit (rx.py:18-19) is materialised: 1 byte per bit
        rc = bbb_clt_run_decimated(c, x0, (first_bit + off) * spb, nb, spb, delay, d_g);
        if (rc) return rc;
        slicer_err_kernel<<<(uint32_t)((nw + 255) / 256), 256, 0, c->stream>>>(d_g, d_tx, nw, 1u, 0u, (int32_t)amplitude,
                                                                                (int32_t)noise_var, d_err);
        c->last_launches++;
        BBB_CUDA(c, cudaGetLastError());
        rc = detect_impl(c, k, first_bit + off, nb, 0, 0, d_err, &d0, ctr);
        if (rc) return rc;
        rounds += c->last_rounds;
    }
    c->last_rounds = rounds;
    c->last_kernel = "prbs_slicer_detect";
    if (det) *det = d0;
    c->nest--;          // close the outer timing region (the guard's decrement is balanced below)
    c->end_call();
    c->nest++;
    return BBB_OK;
}

}  // extern "C"

// ===========================================================================
// Pulse shaper (gateware/bbb/bitshaper.py:12-91) and the full TX word (tx.py:52-81)
// ===========================================================================
// The HDL holds the last 8 PRBS bits in a shift register and, on each of the 8 clocks of a bit
// period, sums 8 ROM words (+c or -c of one 64-tap pulse, selected by the register bit).  As a
// sequence (the relation the reference test pins, bitshaper.py:150-163): y[8m+4] = 2 b[m] - 1, else 0;
// S[t] = sum_k c[k] y[t-k]; the register resets to 0, i.e. b[m] = 0 (a -1 pulse) for m < 0.  The HDL's
// 13-clock pipeline latency is dropped (the same convention as the CLT sample index).
struct ShaperCoef { int16_t c[64]; };

// PRBS bit J (absolute) from words that start at absolute bit `origin`; 0 before the start of time
__device__ __forceinline__ uint32_t tx_bit(const uint64_t *__restrict__ w, int64_t J, int64_t origin)
{
    if (J < 0) return 0u;
    const uint64_t r = (uint64_t)(J - origin);
    return (uint32_t)(w[r >> 6] >> (r & 63)) & 1u;
}

// S[8J+p] for p = 0..7 from the 9 bits b[J-8] .. b[J] (win bit q = b[J-8+q])
__device__ __forceinline__ void shaped8(const ShaperCoef &cf, uint32_t win, int32_t (&s)[8])
{
#pragma unroll
    for (int p = 0; p < 8; p++) {
        int32_t a = 0;
#pragma unroll
        for (int i = 0; i < 8; i++) {
            // p >= 4: tap 8i + p - 4 of bit J - i;  p < 4: tap 8i + p + 4 of bit J - 1 - i
            const int k = p >= 4 ? 8 * i + p - 4 : 8 * i + p + 4;
            const int q = p >= 4 ? 8 - i : 7 - i;
            const int32_t c = cf.c[k];
            a += ((win >> q) & 1u) ? c : -c;
        }
        s[p] = a;
    }
}

// TX words x[t] = wrap12(bit_en * S[t] + g[t] * noise_var), t in [8*first_bit, 8*(first_bit+nbits)); one
// thread per bit period: eight int16 = one 128-bit streaming store.  g may be null (noise off).
// The eight sums of a bit period depend on 8 of the 9 window bits each (phases 0..3 on b[J-8..J-1],
// phases 4..7 on b[J-7..J]), so a CTA first tabulates them: T2[pp][idx] = the sums of phases 2pp and
// 2pp+1 as a packed int16 pair for every 8-bit pattern (8 adds per entry).  A bit period is then 4 LDS.32
// and one store: the kernel is bound by HBM writes, not by the ALU.
#ifndef SHAPER_COPIES
#define SHAPER_COPIES 32                 /* table copies (a power of two <= 32): lane l reads copy l % SHAPER_COPIES */
#endif
#ifndef SHAPER_CTAS
#define SHAPER_CTAS 1
#endif
#ifndef SHAPER_UNROLL
#define SHAPER_UNROLL 4
#endif
#ifndef SHAPER_INTERLEAVE
#define SHAPER_INTERLEAVE 1
#endif
#define SHAPER_THREADS 1024
constexpr int kShaperUnroll = SHAPER_UNROLL;
#define SHAPER_TBYTES (256 * SHAPER_COPIES * 4)      /* one phase-pair table */
#define SHAPER_SMEM (4 * SHAPER_TBYTES)
#define SHAPER_LOG2C (SHAPER_COPIES == 32 ? 5 : SHAPER_COPIES == 16 ? 4 : SHAPER_COPIES == 8 ? 3 : SHAPER_COPIES == 4 ? 2 : SHAPER_COPIES == 2 ? 1 : 0)

// one bit period: table words of the lane's copy (byte offsets), noise add and 12-bit wrap when asked, one 128-bit store
__device__ __forceinline__ void shaper_emit(const char *Tl, uint32_t win, uint64_t r, uint32_t bit_en,
                                            const int8_t *__restrict__ g, int32_t noise_var, uint4 *__restrict__ out)
{
    const uint32_t lo = (win & 255u) << (SHAPER_LOG2C + 2), hi = (win >> 1) << (SHAPER_LOG2C + 2);
    uint32_t o[4] = {*(const uint32_t *)(Tl + lo), *(const uint32_t *)(Tl + SHAPER_TBYTES + lo),
                     *(const uint32_t *)(Tl + 2 * SHAPER_TBYTES + hi), *(const uint32_t *)(Tl + 3 * SHAPER_TBYTES + hi)};
    if (g || !bit_en) {
        uint2 gg = make_uint2(0u, 0u);
        if (g) gg = *(const uint2 *)(g + r * 8);
#pragma unroll
        for (int p = 0; p < 8; p++) {
            const int32_t sp = bit_en ? (int32_t)(int16_t)(o[p >> 1] >> (16 * (p & 1))) : 0;
            const int32_t nz = (int32_t)(int8_t)(((p < 4 ? gg.x : gg.y) >> (8 * (p & 3))) & 0xffu) * noise_var;
            const int32_t x = sp + nz;
            const uint32_t x12 = (uint32_t)((((x + 2048) & 4095) - 2048)) & 0xffffu;
            o[p >> 1] = (p & 1) ? (o[p >> 1] & 0xffffu) | (x12 << 16) : (o[p >> 1] & 0xffff0000u) | x12;
        }
    }
    __stcs(out + r, make_uint4(o[0], o[1], o[2], o[3]));
}

// Table layout: one private copy per lane, word (pp * 256 + idx) * 32 + lane (128 KB): the four lookups of
// a bit period are conflict-free whatever the data (a single 4 KB table costs ~3.5 wavefronts per lookup
// on random patterns, which made shared memory, not HBM, the limit: 4.6 TB/s).  One 1024-thread CTA per SM.
__global__ void __launch_bounds__(SHAPER_THREADS, SHAPER_CTAS)
shaper_stream_kernel(const uint64_t *__restrict__ txw, int64_t origin, int64_t first_bit, uint64_t nbits,
                     const int16_t *__restrict__ cfd, uint32_t bit_en, const int8_t *__restrict__ g, int32_t noise_var,
                     uint4 *__restrict__ out)
{
    extern __shared__ uint32_t T2[];
    __shared__ uint32_t stage[1024];
    __shared__ int16_t cfs[64];           // the 64 taps: indexed at run time, so not a by-value parameter (local memory)
    if (threadIdx.x < 64) cfs[threadIdx.x] = cfd[threadIdx.x];
    __syncthreads();
    {
        const uint32_t e = threadIdx.x, pp = e >> 8, idx = e & 255u;
        uint32_t pair = 0;
#pragma unroll
        for (int h = 0; h < 2; h++) {
            const int p = 2 * (int)pp + h;
            int32_t a = 0;
#pragma unroll
            for (int i = 0; i < 8; i++) {
                // p >= 4: tap 8i + p - 4 of bit J - i;  p < 4: tap 8i + p + 4 of bit J - 1 - i; pattern bit 7 - i
                const int k = p >= 4 ? 8 * i + p - 4 : 8 * i + p + 4;
                const int32_t c = cfs[k];
                a += ((idx >> (7 - i)) & 1u) ? c : -c;
            }
            pair |= ((uint32_t)a & 0xffffu) << (16 * h);
        }
        stage[e] = pair;
    }
    __syncthreads();
    for (uint32_t i = threadIdx.x; i < 4 * 256 * SHAPER_COPIES; i += SHAPER_THREADS) T2[i] = stage[i >> SHAPER_LOG2C];
    __syncthreads();
    // A warp owns a contiguous run of 32-period steps; lane l takes period R + l of each step, so a step is one
    // coalesced 512-byte store.  The 9-bit window of lane l starts at stream bit c0 + R + l: with R a multiple of
    // 32 the shift is a per-lane constant and the window is one funnel shift of two consecutive 32-bit words of
    // the packed PRBS stream, both L1 hits.
    const uint32_t lane = threadIdx.x & 31u;
    const char *const Tl = (const char *)(T2 + (lane & (SHAPER_COPIES - 1u)));
    const uint32_t *const t32 = (const uint32_t *)txw;
    const uint64_t nsteps = (nbits + 31) / 32;
    // this warp's steps: s0, s0 + SS, ... < s1.  SHAPER_INTERLEAVE: a CTA owns a contiguous run and its 32 warps
    // interleave (the CTA writes 16 KB contiguous per round: 148 write streams); else every warp owns its own run
    constexpr uint64_t SS = SHAPER_INTERLEAVE ? SHAPER_THREADS / 32 : 1;
    uint64_t s0, s1;
    if (SHAPER_INTERLEAVE) {
        const uint64_t per = (nsteps + gridDim.x - 1) / gridDim.x;
        s0 = (uint64_t)blockIdx.x * per;
        s1 = s0 + per;
        s0 += threadIdx.x >> 5;
    } else {
        const uint64_t nwarps = (uint64_t)gridDim.x * (SHAPER_THREADS / 32);
        const uint64_t per = (nsteps + nwarps - 1) / nwarps;
        s0 = ((uint64_t)blockIdx.x * (SHAPER_THREADS / 32) + (threadIdx.x >> 5)) * per;
        s1 = s0 + per;
    }
    if (s1 > nsteps) s1 = nsteps;
    if (s0 >= s1) return;
    if (first_bit < 8 && s0 == 0) {               // the start of time (b[m] = 0 for m < 0): bit by bit
        const uint64_t r = lane;
        if (r < nbits) {
            const int64_t J = first_bit + (int64_t)r;
            uint32_t win = 0;
#pragma unroll
            for (int q = 0; q < 9; q++) win |= tx_bit(txw, J - 8 + q, origin) << q;
            shaper_emit(Tl, win, r, bit_en, g, noise_var, out);
        }
        s0 += SS;
        if (s0 >= s1) return;
    }
    const uint64_t a0 = (uint64_t)(first_bit - 8 - origin) + s0 * 32 + lane;   // first window bit of this lane
    const uint32_t sh = (uint32_t)(a0 & 31);
    const uint32_t *wp = t32 + (a0 >> 5);
    uint64_t r = s0 * 32 + lane;
    const uint64_t sf = s1 < nbits / 32 ? s1 : nbits / 32;      // steps below sf have all 32 lanes in range
    uint64_t st = s0;
    if (!g && bit_en) {
        uint4 *op = out + r;
#pragma unroll kShaperUnroll
        for (; st < sf; st += SS) {
            const uint32_t win = __funnelshift_r(wp[0], wp[1], sh);
            wp += SS;
            // byte offsets of the lane's table words
            const uint32_t lo = (win & 255u) << (SHAPER_LOG2C + 2), hi = ((win >> 1) & 255u) << (SHAPER_LOG2C + 2);
            __stcs(op, make_uint4(*(const uint32_t *)(Tl + lo), *(const uint32_t *)(Tl + SHAPER_TBYTES + lo),
                                  *(const uint32_t *)(Tl + 2 * SHAPER_TBYTES + hi), *(const uint32_t *)(Tl + 3 * SHAPER_TBYTES + hi)));
            op += 32 * SS;
        }
        r = st * 32 + lane;
    } else {
        for (; st < sf; st += SS) {
            const uint32_t win = __funnelshift_r(wp[0], wp[1], sh) & 0x1ffu;
            wp += SS;
            shaper_emit(Tl, win, r, bit_en, g, noise_var, out);
            r += 32 * SS;
        }
    }
    if (st < s1 && r < nbits)                     // the ragged last step of the range
        shaper_emit(Tl, __funnelshift_r(wp[0], wp[1], sh) & 0x1ffu, r, bit_en, g, noise_var, out);
}

// RX decisions through the shaped channel: bit J is decided on the sample 8*(J+4)+delay (the pulse of
// bit J peaks at tap 32), sliced at zero (rx.py:28); err word = decisions XOR transmitted bits.
// One thread per 64 decisions.  The shaped sample of a bit period depends on 8 of the 9 window bits, so a CTA
// tabulates it for the call's `delay` (256 int32, one copy per lane: conflict-free lookups); a thread then takes
// its 73 window bits from three packed PRBS words and its 64 noise bytes with four 128-bit loads.
__global__ void __launch_bounds__(256)
shaped_slicer_err_kernel(const uint64_t *__restrict__ txw, int64_t origin, int64_t first_bit, uint64_t nwords,
                         const int16_t *__restrict__ cfd, uint32_t delay, const int8_t *__restrict__ g, int32_t noise_var,
                         uint64_t *__restrict__ err)
{
    __shared__ int32_t S[256 * 32];
    __shared__ int16_t cfs[64];           // run-time indexed taps (a by-value struct would be copied to local memory)
    if (threadIdx.x < 64) cfs[threadIdx.x] = cfd[threadIdx.x];
    __syncthreads();
    {
        const uint32_t idx = threadIdx.x;         // 256 threads: one pattern each
        int32_t a = 0;
#pragma unroll
        for (int t = 0; t < 8; t++) {
            // delay >= 4: tap 8t + delay - 4 of bit J - t (pattern bit 7 - t of win >> 1); else tap 8t + delay + 4 of bit J - 1 - t
            const int k = delay >= 4 ? 8 * t + (int)delay - 4 : 8 * t + (int)delay + 4;
            const int32_t c = cfs[k];
            a += ((idx >> (7 - t)) & 1u) ? c : -c;
        }
#pragma unroll
        for (int l = 0; l < 32; l++) S[idx * 32 + ((l + idx) & 31)] = a;
    }
    __syncthreads();
    const uint64_t j = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= nwords) return;
    const int32_t *const Sl = S + (threadIdx.x & 31u);
    const uint32_t wsh = delay >= 4 ? 1u : 0u;
    uint64_t e = 0;
    if (first_bit + (int64_t)(j * 64) < 4 || ((uintptr_t)g & 15)) {       // the start of time, bit by bit
        for (uint32_t i = 0; i < 64; i++) {
            const uint64_t r = j * 64 + i;
            const int64_t J = first_bit + (int64_t)r + 4;          // the bit period the sample lies in
            uint32_t win = 0;
#pragma unroll
            for (int q = 0; q < 9; q++) win |= tx_bit(txw, J - 8 + q, origin) << q;
            const int32_t x = Sl[((win >> wsh) & 255u) * 32] + (int32_t)g[r] * noise_var;
            const int32_t x12 = ((x + 2048) & 4095) - 2048;
            e |= (uint64_t)((uint32_t)(x12 >= 0) ^ ((win >> 4) & 1u)) << i;   // b[J-4] = the bit being decided
        }
        err[j] = e;
        return;
    }
    const uint64_t a0 = (uint64_t)(first_bit - 4 - origin) + j * 64;       // first window bit: b[J-8] of decision 0
    const uint32_t sh = (uint32_t)(a0 & 63);
    const uint64_t *wp = txw + (a0 >> 6);
    const uint64_t w0 = wp[0], w1 = wp[1], w2 = sh > 55 ? wp[2] : 0ull;
    const uint64_t B0 = sh ? (w0 >> sh) | (w1 << (64 - sh)) : w0;          // window bits 0..63
    const uint64_t B1 = sh ? (w1 >> sh) | (w2 << (64 - sh)) : w1;          // window bits 64..
    const uint4 *gp = (const uint4 *)(g + j * 64);
#pragma unroll
    for (uint32_t q = 0; q < 4; q++) {
        const uint32_t V = (uint32_t)(q ? (B0 >> (16 * q)) | (B1 << (64 - 16 * q)) : B0);   // window bits 16q .. 16q+31
        const uint4 v = __ldcs(gp + q);
        const uint32_t vv[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
        for (uint32_t i = 0; i < 16; i++) {
            const uint32_t win = V >> i;
            const int32_t nz = (int32_t)(int8_t)((vv[i >> 2] >> (8 * (i & 3))) & 0xffu) * noise_var;
            const int32_t x = Sl[((win >> wsh) & 255u) * 32] + nz;
            const int32_t x12 = ((x + 2048) & 4095) - 2048;
            e |= (uint64_t)((uint32_t)(x12 >= 0) ^ ((win >> 4) & 1u)) << (16 * q + i);
        }
    }
    err[j] = e;
}

// PRBS words covering absolute bits [origin, last) on the device (scratch slot 10); origin = 64 bits
// before `first` when there are that many
// Pulser (tx.py:20-31, clocked in the shaper's prbsclk domain, bitshaper.py:35): source bit m is 1 when m % 256 == 0
__global__ void pulse_words_kernel(uint64_t *__restrict__ out, uint64_t org, uint64_t nwords)
{
    for (uint64_t j = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; j < nwords; j += (uint64_t)gridDim.x * blockDim.x) {
        const uint32_t b0 = (uint32_t)((256u - (uint32_t)((org + 64 * j) & 255u)) & 255u);
        out[j] = b0 < 64u ? 1ull << b0 : 0ull;
    }
}

static int shaper_tx_words(bbb_ctx *c, uint32_t k, uint64_t first, uint64_t nbits_with_lookahead, size_t extra_words,
                           uint64_t **d_tx, int64_t *origin)
{
    const uint64_t org = first >= 64 ? first - 64 : 0;
    const uint64_t span = ((first - org) + nbits_with_lookahead + 63) & ~63ull;
    uint64_t *d = (uint64_t *)c->scratch(10, (size_t)(span / 64 + extra_words + 1) * 8);   // + the word a window load may touch
    if (!d) return BBB_ECUDA;
    int rc = BBB_OK;
    if (k == BBB_TX_SRC_PULSE) {
        const uint64_t nwords = span / 64;
        const uint64_t blocks = (nwords + 255) / 256, most = (uint64_t)c->sm_count * 8;
        pulse_words_kernel<<<(uint32_t)(blocks < most ? (blocks ? blocks : 1) : most), 256, 0, c->stream>>>(d, org, nwords);
        c->last_launches++;
        BBB_CUDA(c, cudaGetLastError());
    } else {
        rc = bbb_prbs_stream_dev(c, k, 1u, org, span, d);
    }
    if (rc) return rc;
    *d_tx = d;
    *origin = (int64_t)org;
    return BBB_OK;
}

static int shaper_check(bbb_ctx *c, const int16_t *coeff, ShaperCoef &cf)
{
    if (!coeff) return c->fail(BBB_EINVAL, "coeff is null");
    for (int i = 0; i < 64; i++) {
        if (coeff[i] <= -256 || coeff[i] > 255) return c->fail(BBB_EINVAL, "coefficients must lie in (-256, 255] (bitshaper.py:19-20)");
        cf.c[i] = coeff[i];
    }
    return BBB_OK;
}

// the 64 taps on the device (scratch slot 14): the kernels index them at run time
static const int16_t *shaper_upload(bbb_ctx *c, const ShaperCoef &cf)
{
    int16_t *d = (int16_t *)c->scratch(14, sizeof cf.c);
    if (!d) return nullptr;
    if (c->h2d(d, cf.c, sizeof cf.c) != cudaSuccess) { c->cuda_fail(cudaGetLastError(), "upload shaper taps"); return nullptr; }
    return d;
}

extern "C" {

int bbb_tx_stream_dev(bbb_ctx *c, uint32_t k, uint64_t first_bit, uint64_t nbits, const int16_t *coeff, uint32_t bit_en,
                      const uint64_t *x0, uint32_t noise_var, int16_t *out_dev)
{
    if (!c) return BBB_EINVAL;
    int rc = k == BBB_TX_SRC_PULSE ? BBB_OK : check_prbs_args(c, k, 0);
    if (rc) return rc;
    BBB_CUDA(c, cudaSetDevice(c->device));
    ShaperCoef cf;
    rc = shaper_check(c, coeff, cf);
    if (rc) return rc;
    if (!out_dev || ((uintptr_t)out_dev & 15)) return c->fail(BBB_EINVAL, "out_dev must be 16-byte aligned");
    if (noise_var > 15) return c->fail(BBB_EINVAL, "noise_var 0..15 (tx.py:49)");
    if (x0 && (c->n == 0 || c->n > 256)) return c->fail(BBB_ESTATE, "load a recurrence with n <= 256 first");
    if ((first_bit + nbits) >> 58) return c->fail(BBB_EINVAL, "bit range too large");
    c->begin_call();
    struct Nest { bbb_ctx *c; Nest(bbb_ctx *cc) : c(cc) { c->nest++; } ~Nest() { c->nest--; } } nest(c);
    const int16_t *d_cf = shaper_upload(c, cf);
    if (!d_cf) return BBB_ECUDA;
    // noise on: 2^29 bits per pass bounds the int8 sample buffer at 4 GiB; noise off: the packed PRBS words
    // are the only intermediate (1/128 of the output), 2^32 bits per pass
    const uint64_t CH = c->tune_slicer_chunk_bits ? c->tune_slicer_chunk_bits : (x0 ? (1ull << 29) : (1ull << 32));
    for (uint64_t off = 0; off < nbits; off += CH) {
        const uint64_t nb = nbits - off < CH ? nbits - off : CH;
        uint64_t *d_tx;
        int64_t origin;
        rc = shaper_tx_words(c, k, first_bit + off, nb, 0, &d_tx, &origin);
        if (rc) return rc;
        int8_t *d_g = nullptr;
        if (x0) {
            d_g = (int8_t *)c->scratch(9, (size_t)nb * 8);
            if (!d_g) return BBB_ECUDA;
            rc = bbb_clt_stream_dev(c, x0, (first_bit + off) * 8, nb * 8, BBB_MODEL_GATEWARE, 1u, d_g, nullptr);
            if (rc) return rc;
        }
        const uint64_t want_blocks = (nb + SHAPER_THREADS - 1) / SHAPER_THREADS, max_blocks = (uint64_t)c->sm_count * SHAPER_CTAS;
        BBB_CUDA(c, cudaFuncSetAttribute(shaper_stream_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, SHAPER_SMEM));
        shaper_stream_kernel<<<(uint32_t)(want_blocks < max_blocks ? want_blocks : max_blocks), SHAPER_THREADS, SHAPER_SMEM, c->stream>>>(
            d_tx, origin, (int64_t)(first_bit + off), nb, d_cf, bit_en, d_g, (int32_t)noise_var, (uint4 *)(out_dev + off * 8));
        c->last_launches++;
        BBB_CUDA(c, cudaGetLastError());
    }
    c->last_kernel = "tx_stream";
    c->nest--;
    c->end_call();
    c->nest++;
    return BBB_OK;
}

int bbb_tx_stream(bbb_ctx *c, uint32_t k, uint64_t first_bit, uint64_t nbits, const int16_t *coeff, uint32_t bit_en,
                  const uint64_t *x0, uint32_t noise_var, int16_t *out)
{
    if (!c) return BBB_EINVAL;
    if (!out) return c->fail(BBB_EINVAL, "out is null");
    return bbb_stream_to_host(c, 12, out, (size_t)nbits * 16, 16 * 4096, (size_t)(x0 ? 256 : 64) << 20,
                              [&](size_t off, size_t bytes, void *dev) {
                                  return bbb_tx_stream_dev(c, k, first_bit + off / 16, bytes / 16, coeff, bit_en, x0,
                                                           noise_var, (int16_t *)dev);
                              });
}

int bbb_prbs_shaped_slicer_detect(bbb_ctx *c, uint32_t k, uint64_t first_bit, uint64_t nbits, const int16_t *coeff,
                                  const uint64_t *x0, uint32_t noise_var, uint32_t delay, bbb_det_state *det,
                                  bbb_det_counters *ctr)
{
    int rc = check_prbs_args(c, k, 0);
    if (rc) return rc;
    ShaperCoef cf;
    rc = shaper_check(c, coeff, cf);
    if (rc) return rc;
    if (!ctr || !x0) return c->fail(BBB_EINVAL, "null argument");
    if (nbits & 63) return c->fail(BBB_EINVAL, "nbits must be a multiple of 64");
    if (noise_var > 15 || delay > 7) return c->fail(BBB_EINVAL, "noise_var 0..15, delay 0..7");
    if (c->n == 0 || c->n > 256) return c->fail(BBB_ESTATE, "load a recurrence with n <= 256 first");
    if ((first_bit + nbits) >> 58) return c->fail(BBB_EINVAL, "bit range too large");
    bbb_det_state d0;
    if (det) d0 = *det;
    else { d0.bit_in = 0; d0.prbs = 1; d0.err_sr = (uint32_t)((1ull << k) - 1); d0.reload_ctr = 0; }
    c->begin_call();
    struct Nest { bbb_ctx *c; Nest(bbb_ctx *cc) : c(cc) { c->nest++; } ~Nest() { c->nest--; } } nest(c);
    const uint64_t CH = c->tune_slicer_chunk_bits ? c->tune_slicer_chunk_bits : (1ull << 32);   // 4 GiB of sampled int8 noise per pass
    const int16_t *d_cf = shaper_upload(c, cf);
    if (!d_cf) return BBB_ECUDA;
    uint32_t rounds = 0;
    for (uint64_t off = 0; off < nbits; off += CH) {
        const uint64_t nb = nbits - off < CH ? nbits - off : CH, nw = nb / 64;
        uint64_t *d_tx;
        int64_t origin;
        rc = shaper_tx_words(c, k, first_bit + off, nb + 64, nw, &d_tx, &origin);   // + the 4 bits after the range
        if (rc) return rc;
        const uint64_t span_words = (((first_bit + off) - (uint64_t)origin) + nb + 64 + 63) / 64;
        uint64_t *d_err = d_tx + span_words;
        int8_t *d_g = (int8_t *)c->scratch(9, (size_t)nb);
        if (!d_g) return BBB_ECUDA;
        // the sample of bit J is 8*(J+4)+delay: only that one of the 8 noise samples is materialised
        rc = bbb_clt_run_decimated(c, x0, (first_bit + off + 4) * 8, nb, 8u, delay, d_g);
        if (rc) return rc;
        shaped_slicer_err_kernel<<<(uint32_t)((nw + 255) / 256), 256, 0, c->stream>>>(
            d_tx, origin, (int64_t)(first_bit + off), nw, d_cf, delay, d_g, (int32_t)noise_var, d_err);
        c->last_launches++;
        BBB_CUDA(c, cudaGetLastError());
        rc = detect_impl(c, k, first_bit + off, nb, 0, 0, d_err, &d0, ctr);
        if (rc) return rc;
        rounds += c->last_rounds;
    }
    c->last_rounds = rounds;
    c->last_kernel = "prbs_shaped_slicer_detect";
    if (det) *det = d0;
    c->nest--;
    c->end_call();
    c->nest++;
    return BBB_OK;
}

}  // extern "C"
