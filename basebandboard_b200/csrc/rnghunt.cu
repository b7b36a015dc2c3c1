// rnghunt.cu -- the rnghunt worker loop (software/rnghunt/src/bin/rnghunt.rs:23-57) on the device.
//
// Per candidate: BinaryMatrix::random (binary_matrix.rs:81-101) -> recur 2n steps from the all-ones
// vector (rnghunt.rs:28-29, binary_matrix.rs:68-76) -> Berlekamp-Massey of the (reversed) output
// sequence (rnghunt.rs:31-36, berlekamp_massey.rs:5-31) -> degree == n (rnghunt.rs:39-41) ->
// is_primitive (binary_polynomial.rs:178-212).
//
// Device pipeline (DESIGN.md "rnghunt path"):
//   1. hunt_front_kernel, one THREAD per candidate: generates the matrix as 4 tap slots per row
//      (row weights are 3 or 4, rnghunt.rs:25) with the greedy column balancing done on bit masks
//      in registers; the recurrence state lives in shared memory as one byte per state bit
//      (word [bit][thread], the 4 bytes of a word are 4 successive generations, so a gather is one
//      conflict-free LDS.U8); every step's output bit feeds an online Berlekamp-Massey, so the 2n
//      sequence bits are never stored.  Candidates of degree n with constant term 1 and an odd
//      number of terms are appended to a survivor list (stream compaction by atomic counter).
//   2. hunt_check_kernel, one thread per (survivor, exponent): x^e mod p by square-and-multiply-by-x;
//      first e = 2^n - 1 for every survivor, then the cofactors (2^n-1)/q for those that passed.
//   3. hunt_classify_kernel: stage per candidate, accepted candidates appended to the hit list.
// The Berlekamp-Massey polynomial of a sequence of length 2n with linear complexity <= n is unique,
// so the online formulation returns exactly the reference's polynomial (parity-tested through
// bbb_rnghunt_eval's poly_words).
#include <algorithm>

#include "bbb_internal.h"
#include "mersenne_factors.inc"

#define HUNT_PENDING 0xFFu

// ---------------------------------------------------------------------------
// candidate stream (DESIGN.md "rnghunt determinism"): draw #i of candidate (seed, c)
// ---------------------------------------------------------------------------
__host__ __device__ __forceinline__ uint64_t hunt_mix64(uint64_t z)
{
    z ^= z >> 30; z *= 0xBF58476D1CE4E5B9ull;
    z ^= z >> 27; z *= 0x94D049BB133111EBull;
    z ^= z >> 31;
    return z;
}

struct HuntRng {
    uint64_t base, i;
    __device__ __forceinline__ void init(uint64_t seed, uint64_t c)
    {
        base = hunt_mix64(seed * 0x9E3779B97F4A7C15ull + c);
        i = 0;
    }
    // uniform-ish index in [0, m)
    __device__ __forceinline__ uint32_t draw(uint32_t m)
    {
        i += 1;
        const uint64_t z = hunt_mix64(base + i * 0x9E3779B97F4A7C15ull);
        return (uint32_t)(((z >> 32) * (uint64_t)m) >> 32);
    }
};

// position of the (r+1)-th set bit of w (r < popc(w))
__device__ __forceinline__ uint32_t nth_set_bit(uint32_t w, uint32_t r)
{
    uint32_t pos = 0;
#pragma unroll
    for (int s = 16; s; s >>= 1) {
        const uint32_t lo = w & ((1u << s) - 1u);
        const uint32_t pc = __popc(lo);
        if (r >= pc) { r -= pc; w >>= s; pos += s; }
        else w = lo;
    }
    return pos;
}

// BinaryMatrix::random(n, n, [3,4,4,4,4,4,4,4]) as tap lists.  All column weights are either the
// current minimum or minimum + 1 (picks only ever raise a minimum-weight column), so the greedy rule
// "uniform among the minimum-weight columns not yet used in this row" (binary_matrix.rs:88-94) runs
// on three bit masks.  put(row, slot, col) receives the picks; slot 3 of a weight-3 row gets `n`.
template <int NW, typename Put>
__device__ __forceinline__ void hunt_generate(uint32_t n, uint64_t seed, uint64_t cand, Put put)
{
    HuntRng rng;
    rng.init(seed, cand);
    uint32_t minm[NW], nextm[NW], rowm[NW];
#pragma unroll
    for (int w = 0; w < NW; w++) {
        const uint32_t lo = (uint32_t)w * 32u;
        minm[w] = n >= lo + 32 ? 0xffffffffu : (n > lo ? (1u << (n - lo)) - 1u : 0u);
        nextm[w] = 0;
    }
    for (uint32_t row = 0; row < n; row++) {
        const uint32_t weight = rng.draw(8) == 0 ? 3u : 4u;   // rowweights[..] = [3,4,4,4,4,4,4,4]
#pragma unroll
        for (int w = 0; w < NW; w++) rowm[w] = 0;
        for (uint32_t k = 0; k < 4; k++) {
            if (k >= weight) { put(row, k, n); continue; }
            uint32_t total = 0;
#pragma unroll
            for (int w = 0; w < NW; w++) total += __popc(minm[w] & ~rowm[w]);
            if (total == 0) { put(row, k, n); continue; }   // the reference would panic (:93 unwrap)
            uint32_t r = rng.draw(total);
            uint32_t word = 0, wi = 0;
            bool found = false;
#pragma unroll
            for (int w = 0; w < NW; w++) {
                const uint32_t cw = minm[w] & ~rowm[w];
                const uint32_t pc = __popc(cw);
                if (!found) {
                    if (r < pc) { found = true; word = cw; wi = (uint32_t)w; }
                    else r -= pc;
                }
            }
            const uint32_t bit = nth_set_bit(word, r);
            const uint32_t m = 1u << bit;
            put(row, k, wi * 32u + bit);
            uint32_t any = 0;
#pragma unroll
            for (int w = 0; w < NW; w++) {
                const uint32_t mm = (uint32_t)w == wi ? m : 0u;
                minm[w] &= ~mm;
                nextm[w] |= mm;
                rowm[w] |= mm;
                any |= minm[w];
            }
            if (!any) {
#pragma unroll
                for (int w = 0; w < NW; w++) { minm[w] = nextm[w]; nextm[w] = 0; }
            }
        }
    }
}

// The same generator with less bookkeeping (n <= 256; used by hunt_front_pow_kernel, where generation is 40 % of the
// instructions).  Within one pass over the columns "raised" = ALL \ minm, so no second mask is kept, and a column
// picked in this row leaves minm at once, so "not yet used in this row" only matters after minm has been refilled
// in the middle of a row (once in ~50 rows): then the row's earlier picks are excluded through `excl`, rebuilt from
// the (at most four) picks.  The number of candidates is counted, not re-derived by popcounts.  Draw for draw the
// same sequence as hunt_generate (the GPU tests compare the kernels that use either with the oracle).
template <int NW, typename Put>
__device__ __forceinline__ void hunt_generate_fast(uint32_t n, uint64_t seed, uint64_t cand, Put put)
{
    HuntRng rng;
    rng.init(seed, cand);
    uint32_t minm[NW], excl[NW];
#pragma unroll
    for (int w = 0; w < NW; w++) {
        const uint32_t lo = (uint32_t)w * 32u;
        minm[w] = n >= lo + 32 ? 0xffffffffu : (n > lo ? (1u << (n - lo)) - 1u : 0u);
        excl[w] = 0;
    }
    uint32_t cntmin = n;
    for (uint32_t row = 0; row < n; row++) {
        const uint32_t weight = rng.draw(8) == 0 ? 3u : 4u;   // rowweights[..] = [3,4,4,4,4,4,4,4]
        uint32_t pb = 0, cols = 0;                            // picks of this row made before a refill; the picks
        for (uint32_t k = 0; k < 4; k++) {
            const uint32_t tot = cntmin - pb;                 // |minm \ row picks|
            if (k >= weight || tot == 0) { put(row, k, n); continue; }
            const uint32_t r = rng.draw(tot);
            uint32_t word = minm[0] & ~excl[0], wi = 0, base = 0, s = __popc(word);
#pragma unroll
            for (int w = 1; w < NW; w++) {
                const uint32_t a = minm[w] & ~excl[w];
                const bool ge = r >= s;
                wi = ge ? (uint32_t)w : wi;
                base = ge ? s : base;
                word = ge ? a : word;
                s += __popc(a);
            }
            const uint32_t bit = nth_set_bit(word, r - base);
            const uint32_t col = wi * 32u + bit, m = 1u << bit;
            put(row, k, col);
#pragma unroll
            for (int w = 0; w < NW; w++) minm[w] &= ~((uint32_t)w == wi ? m : 0u);
            cols |= col << (8 * k);
            if (--cntmin == 0) {                              // every column has been raised: the next level starts
#pragma unroll
                for (int w = 0; w < NW; w++) {
                    const uint32_t lo = (uint32_t)w * 32u;
                    minm[w] = n >= lo + 32 ? 0xffffffffu : (n > lo ? (1u << (n - lo)) - 1u : 0u);
                    excl[w] = 0;
                }
                cntmin = n;
                pb = k + 1;
                for (uint32_t q = 0; q <= k; q++) {
                    const uint32_t c = (cols >> (8 * q)) & 0xffu;
#pragma unroll
                    for (int w = 0; w < NW; w++) excl[w] |= (c >> 5) == (uint32_t)w ? 1u << (c & 31u) : 0u;
                }
            }
        }
        if (pb) {
#pragma unroll
            for (int w = 0; w < NW; w++) excl[w] = 0;
        }
    }
}

// ---------------------------------------------------------------------------
// online Berlekamp-Massey over GF(2) (berlekamp_massey.rs:5-31 fed u[t] = recur output t)
//   c, b: connection polynomials, bit j = coefficient c_j (LSB-first words); bsh = b * x^(t-m)
//   win: bit j = u[t-j]
// NWP = ceil((n+1)/32): a sequence produced by an n-dimensional recurrence has complexity <= n.
// ---------------------------------------------------------------------------
template <int NWP>
struct HuntBM {
    uint32_t c[NWP], bsh[NWP], win[NWP];
    uint32_t l, t;
    __device__ __forceinline__ void init()
    {
#pragma unroll
        for (int w = 0; w < NWP; w++) { c[w] = 0; bsh[w] = 0; win[w] = 0; }
        c[0] = 1;
        bsh[0] = 1;
        l = 0;
        t = 0;
    }
    __device__ __forceinline__ void push(uint32_t bit)
    {
        // slide the window and b * x^(t-m) by one position
#pragma unroll
        for (int w = NWP - 1; w > 0; w--) {
            win[w] = (win[w] << 1) | (win[w - 1] >> 31);
            bsh[w] = (bsh[w] << 1) | (bsh[w - 1] >> 31);
        }
        win[0] = (win[0] << 1) | (bit & 1u);
        bsh[0] <<= 1;
        uint32_t d = 0;
#pragma unroll
        for (int w = 0; w < NWP; w++) d ^= c[w] & win[w];
        if (__popc(d) & 1) {                                   // discrepancy (:17)
            const bool grow = 2 * l <= t;                      // l <= n/2 (:23)
#pragma unroll
            for (int w = 0; w < NWP; w++) {
                const uint32_t tc = c[w];
                c[w] ^= bsh[w];                                // c ^= b >> (n - m) (:21-22)
                if (grow) bsh[w] = tc;
            }
            if (grow) l = t + 1 - l;
        }
        t++;
    }
};

// ---------------------------------------------------------------------------
// outputs of the front kernels
// ---------------------------------------------------------------------------
struct HuntOut {
    uint8_t *stage;            // [count] BBB_STAGE_DEGREE / _FILTER / HUNT_PENDING
    uint32_t *surv_count;      // survivors appended so far
    uint32_t *surv_cand;       // [cap] candidate slot (index into stage[])
    uint32_t *surv_poly;       // [cap][NWP] p(x), bit i = coefficient of x^i
    uint64_t *poly_out;        // [count][ceil((n+1)/64)] reference layout, or null
};

// a[idx] for a register array and a run-time index (0 when idx is out of range): a select chain, so the
// array never goes to local memory
template <int N>
__device__ __forceinline__ uint32_t reg_pick(const uint32_t (&a)[N], uint32_t idx)
{
    uint32_t r = 0;
#pragma unroll
    for (int k = 0; k < N; k++) r = (uint32_t)k == idx ? a[k] : r;
    return r;
}

// Degree / filter verdict of one candidate and, for survivors, its polynomial.
// c: connection polynomial bits c_j (the reference's coefficient j, highest power first).
template <int NWP>
__device__ __forceinline__ void hunt_finish(const HuntOut &o, uint32_t n, uint32_t slot, const uint32_t (&c)[NWP],
                                         uint32_t l)
{
    const uint32_t pw64 = (n + 1 + 63) / 64;
    if (o.poly_out) {
        // BinaryPolynomial words: coefficient j at word j/64, bit 63 - j%64 (binary_vector.rs:25-31)
#pragma unroll
        for (int w = 0; w < (NWP + 1) / 2; w++) {
            if ((uint32_t)w < pw64) {
                uint64_t v = 0;
                if (l == n) {
                    const uint32_t lo = c[2 * w];
                    const uint32_t hi = 2 * w + 1 < NWP ? c[2 * w + 1 < NWP ? 2 * w + 1 : 0] : 0u;
                    v = __brevll(((uint64_t)hi << 32) | lo);
                }
                o.poly_out[(size_t)slot * pw64 + w] = v;
            }
        }
    }
    if (l != n) { o.stage[slot] = BBB_STAGE_DEGREE; return; }               // rnghunt.rs:39-41
    uint32_t ones = 0;
#pragma unroll
    for (int w = 0; w < NWP; w++) ones += __popc(c[w]);
    const uint32_t cn = (reg_pick<NWP>(c, n >> 5) >> (n & 31)) & 1u;        // constant term = c_n
    if (!cn || !(ones & 1u)) { o.stage[slot] = BBB_STAGE_FILTER; return; }  // binary_polynomial.rs:185-192
    o.stage[slot] = HUNT_PENDING;
    const uint32_t s = atomicAdd(o.surv_count, 1u);
    o.surv_cand[s] = slot;
    // p(x) = sum c_j x^(n-j): reverse the whole register array, then shift the n+1 coefficients down
    uint32_t rev[NWP];
#pragma unroll
    for (int w = 0; w < NWP; w++) rev[w] = __brev(c[NWP - 1 - w]);
    const uint32_t sh = 32u * NWP - 1u - n, ws = sh >> 5, bs = sh & 31u;
    uint32_t *dst = o.surv_poly + (size_t)s * NWP;
#pragma unroll
    for (int w = 0; w < NWP; w++)
        dst[w] = __funnelshift_r(reg_pick<NWP>(rev, (uint32_t)w + ws), reg_pick<NWP>(rev, (uint32_t)w + ws + 1u), bs);
}

__device__ __forceinline__ uint32_t lds_u8(uint32_t sa)
{
    uint32_t v;
    asm volatile("ld.shared.u8 %0, [%1];" : "=r"(v) : "r"(sa));
    return v;
}
__device__ __forceinline__ uint32_t lds_u32(uint32_t sa)
{
    uint32_t v;
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(sa));
    return v;
}
// byte k of t (sel = 0x4440 + k); volatile so that the compiler does not hoist the 4n step-invariant
// extractions out of the step loop and spill them
__device__ __forceinline__ uint32_t tap_byte(uint32_t t, uint32_t sel)
{
    uint32_t d;
    asm volatile("prmt.b32 %0, %1, %2, %3;" : "=r"(d) : "r"(t), "r"(0u), "r"(sel));
    return d;
}
__device__ __forceinline__ void sts_u8(uint32_t sa, uint32_t v)
{
    asm volatile("st.shared.u8 [%0], %1;" ::"r"(sa), "r"(v) : "memory");
}

// ---------------------------------------------------------------------------
// front kernel for generated candidates: one thread per candidate
// ---------------------------------------------------------------------------
struct FrontParams {
    uint32_t n;
    uint64_t seed, first, stride;
    uint32_t count;
    HuntOut out;
    uint16_t *dump_taps;   // [n][4] taps of candidate `first` (bbb_rnghunt_candidate), else null
    uint32_t *tap_scratch; // hunt_front_pow_kernel: [grid][n][32] words, private to the launch
    uint32_t *group_counter;   // hunt_front_pow_kernel: zeroed before the launch
};

// shared memory: state bytes xs[(n+1)][B] words, then taps[n][B] (4 x TapT per thread and row)
template <int NW, typename TapT>
__global__ void hunt_front_kernel(FrontParams p)
{
    constexpr int NWP = NW + 1;   // n + 1 bits always fit
    extern __shared__ uint32_t hsm[];
    const uint32_t B = blockDim.x, tid = threadIdx.x, n = p.n;
    const uint32_t slot = blockIdx.x * B + tid;
    uint32_t *xs = hsm;                                   // word (bit, thread): 4 generations
    TapT *taps = (TapT *)(hsm + (size_t)(n + 1) * B);     // [row][thread][4]
    if (slot >= p.count) return;
    const uint64_t cand = p.first + (uint64_t)slot * p.stride;

    hunt_generate<NW>(n, p.seed, cand, [&](uint32_t row, uint32_t k, uint32_t col) {
        taps[((size_t)row * B + tid) * 4 + k] = (TapT)col;
        if (p.dump_taps && slot == 0) p.dump_taps[row * 4 + k] = (uint16_t)(col == n ? 0xFFFFu : col);
    });
    if (p.dump_taps) return;

    // b = all ones (rnghunt.rs:28): generation 0; the extra row n stays 0 (empty tap slots)
    for (uint32_t i = 0; i < n; i++) xs[i * B + tid] = 1u;
    xs[n * B + tid] = 0u;

    HuntBM<NWP> bm;
    bm.init();
    const uint32_t rs = 4 * B;                            // byte stride between state bits
    constexpr uint32_t RB = 8;                            // rows per batch: all loads, then all stores
    const uint32_t nfull = n / RB * RB;
    const uint32_t xs_sa = (uint32_t)__cvta_generic_to_shared(xs + tid);      // shared-window byte addresses
    const uint32_t tp_sa = (uint32_t)__cvta_generic_to_shared(taps + (size_t)tid * 4);
    const uint32_t tps = B * 4 * (uint32_t)sizeof(TapT);                       // bytes between rows of taps
    for (uint32_t step = 0; step < 2 * n; step++) {
        const uint32_t xo = xs_sa + (step & 3), xn = xs_sa + ((step + 1) & 3);
        // x' = A . x (binary_matrix.rs:53-64 / :72): row r gathers its 3-4 taps.  The gathers of a
        // batch are issued before any store so that they overlap; explicit shared-memory addresses
        // keep a gather at three instructions (byte extract, multiply-add, LDS.U8).
        for (uint32_t r0 = 0; r0 < nfull; r0 += RB) {
            uint32_t v[RB];
            if (sizeof(TapT) == 1) {
                uint32_t t4[RB];
#pragma unroll
                for (uint32_t k = 0; k < RB; k++) t4[k] = lds_u32(tp_sa + (r0 + k) * tps);
#pragma unroll
                for (uint32_t k = 0; k < RB; k++)
                    v[k] = lds_u8(__byte_perm(t4[k], 0, 0x4440) * rs + xo) ^ lds_u8(__byte_perm(t4[k], 0, 0x4441) * rs + xo) ^
                           lds_u8(__byte_perm(t4[k], 0, 0x4442) * rs + xo) ^ lds_u8(__byte_perm(t4[k], 0, 0x4443) * rs + xo);
            } else {
                uint32_t ta[RB], tb[RB];
#pragma unroll
                for (uint32_t k = 0; k < RB; k++) {
                    ta[k] = lds_u32(tp_sa + (r0 + k) * tps);
                    tb[k] = lds_u32(tp_sa + (r0 + k) * tps + 4);
                }
#pragma unroll
                for (uint32_t k = 0; k < RB; k++)
                    v[k] = lds_u8((ta[k] & 0xffffu) * rs + xo) ^ lds_u8((ta[k] >> 16) * rs + xo) ^
                           lds_u8((tb[k] & 0xffffu) * rs + xo) ^ lds_u8((tb[k] >> 16) * rs + xo);
            }
#pragma unroll
            for (uint32_t k = 0; k < RB; k++) sts_u8(xn + (r0 + k) * rs, v[k]);
        }
        for (uint32_t r = nfull; r < n; r++) {
            uint32_t c[4];
#pragma unroll
            for (int k = 0; k < 4; k++) c[k] = taps[((size_t)r * B + tid) * 4 + k];
            const uint32_t v = lds_u8(c[0] * rs + xo) ^ lds_u8(c[1] * rs + xo) ^ lds_u8(c[2] * rs + xo) ^ lds_u8(c[3] * rs + xo);
            sts_u8(xn + r * rs, v);
        }
        bm.push(lds_u8(xn));                              // x[0] after the step (binary_matrix.rs:73)
    }
    hunt_finish<NWP>(p.out, n, slot, bm.c, bm.l);
}

// ---------------------------------------------------------------------------
// front kernel, one WARP per candidate (n <= 256): the product path.
//
// The state x (n bits = NW words) is replicated in every lane; lane l owns rows l, l+32, ... and keeps
// each of them as NW mask words (bit c of word w set = column 32w+c is a tap of the row), so a row is
// NW LOP3 ((W & M) ^ acc) + one parity, and the NW ballots of a step ARE the new state words (lane l's
// vote in ballot j = row l+32j): no gather, no shared-memory traffic and no per-step address arithmetic.
// A step costs NW*NW + 3*NW warp instructions for all n rows, where the thread-per-candidate kernel
// spends 16 instructions per row and thread (4 LDS.U8 gathers + 1 STS.U8 per row, shared-memory bound).
//
// BinaryMatrix::random and Berlekamp-Massey are sequential per candidate, so they run thread-per-
// candidate around the warp phase: CPW lanes generate CPW candidates (taps go to shared memory, one
// word per row, an empty fourth slot repeating the third pick), the warp then clocks the candidates one
// after the other and leaves each one's 2n output bits in shared memory, and finally lane c runs the
// online Berlekamp-Massey over candidate c's bits.  Nothing but the verdict (and the polynomial of the
// few survivors) goes to global memory.
// shared: taps[n][CPW+1] words, seq[ceil(2n/32)][CPW] words
// ---------------------------------------------------------------------------
// The per-warp tap scratch is rewritten for every group of candidates: keep its lines in L2 (evict_last) so that
// they are not written back to HBM between two uses.
__device__ __forceinline__ uint64_t l2_keep_policy()
{
    uint64_t pol;
    asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(pol));
    return pol;
}
__device__ __forceinline__ void st_keep(uint32_t *p, uint32_t v, uint64_t pol)
{
    asm volatile("st.global.L2::cache_hint.u32 [%0], %1, %2;" ::"l"(p), "r"(v), "l"(pol) : "memory");
}
__device__ __forceinline__ uint32_t ld_keep(const uint32_t *p, uint64_t pol)
{
    uint32_t v;
    asm volatile("ld.global.cg.L2::cache_hint.u32 %0, [%1], %2;" : "=r"(v) : "l"(p), "l"(pol) : "memory");
    return v;
}

// ballot of (x & 1) over the warp: the AND and the predicate in ONE LOP3 (its predicate output), then the vote --
// written out because the compiler spends LOP3 + ISETP on `__ballot_sync(~0u, x & 1)`
__device__ __forceinline__ uint32_t ballot_odd(uint32_t x)
{
    uint32_t r;
    asm volatile("{\n\t.reg .pred p, q;\n\t.reg .b32 d;\n\tsetp.ne.u32 q, 0, 0;\n\t"
                 "lop3.or.b32 d|p, %1, 1, 0, 0xc0, q;\n\t"
                 "vote.sync.ballot.b32 %0, p, 0xffffffff;\n\t}"
                 : "=r"(r) : "r"(x));
    return r;
}

template <int NW, int CPW>
__global__ void __launch_bounds__(32, 8)
hunt_front_warp_kernel(FrontParams p)
{
    constexpr int NWP = NW + 1;
    constexpr uint32_t TS = CPW + 1;                       // odd stride: conflict-free both ways
    extern __shared__ uint32_t hsm[];
    const uint32_t lane = threadIdx.x, n = p.n;
    const uint32_t slot0 = blockIdx.x * CPW;
    if (slot0 >= p.count) return;
    const uint32_t ncand = p.count - slot0 < (uint32_t)CPW ? p.count - slot0 : (uint32_t)CPW;
    const uint32_t SW = (2 * n + 31) / 32;                 // sequence words per candidate (<= 2 NW)
    uint32_t *taps = hsm;
    uint32_t *seq = hsm + (size_t)n * TS;

    // phase 1: lane c generates candidate slot0 + c (binary_matrix.rs:81-101)
    if (lane < ncand) {
        const uint64_t cand = p.first + (uint64_t)(slot0 + lane) * p.stride;
        uint32_t acc = 0, prev = 0;
        hunt_generate<NW>(n, p.seed, cand, [&](uint32_t row, uint32_t k, uint32_t col) {
            const uint32_t cc = col == n ? prev : col;     // empty slot: repeat the previous pick (masks are ORed)
            prev = cc;
            acc |= cc << (8 * k);
            if (k == 3) { taps[row * TS + lane] = acc; acc = 0; }
        });
    }
    __syncwarp();

    // phase 2: the warp clocks one candidate at a time (binary_matrix.rs:68-76 from b = all ones, rnghunt.rs:28-29)
    for (uint32_t c = 0; c < ncand; c++) {
        uint32_t M[NW][NW];
#pragma unroll
        for (int j = 0; j < NW; j++) {
#pragma unroll
            for (int w = 0; w < NW; w++) M[j][w] = 0;
            const uint32_t row = lane + 32u * j;
            if (row < n) {
                const uint32_t t4 = taps[row * TS + c];
#pragma unroll
                for (int k = 0; k < 4; k++) {
                    const uint32_t t = (t4 >> (8 * k)) & 0xffu, bit = 1u << (t & 31u);
#pragma unroll
                    for (int w = 0; w < NW; w++) M[j][w] |= (t >> 5) == (uint32_t)w ? bit : 0u;
                }
            }
        }
        uint32_t W[NW];
#pragma unroll
        for (int w = 0; w < NW; w++) {
            const uint32_t lo = (uint32_t)w * 32u;
            W[w] = n >= lo + 32 ? 0xffffffffu : (n > lo ? (1u << (n - lo)) - 1u : 0u);
        }
        uint32_t acc = 0, mine = 0;
        for (uint32_t ch = 0; ch < SW; ch++) {
#pragma unroll 4
            for (int st = 0; st < 32; st++) {
                uint32_t Wn[NW];
#pragma unroll
                for (int j = 0; j < NW; j++) {
                    uint32_t v = W[0] & M[j][0];
#pragma unroll
                    for (int w = 1; w < NW; w++) v ^= W[w] & M[j][w];
                    Wn[j] = ballot_odd(__popc(v));
                }
#pragma unroll
                for (int w = 0; w < NW; w++) W[w] = Wn[w];
                acc = (acc << 1) | (W[0] & 1u);            // x[0] after the step (binary_matrix.rs:73); first bit ends on top
            }
            if (lane == ch) mine = acc;
        }
        if (lane < SW) seq[lane * CPW + c] = mine;
    }
    __syncwarp();

    // phase 3: lane c runs Berlekamp-Massey over candidate c's 2n bits (berlekamp_massey.rs:5-31)
    if (lane < ncand) {
        HuntBM<NWP> bm;
        bm.init();
        for (uint32_t ch = 0; ch < SW; ch++) {
            const uint32_t word = seq[ch * CPW + lane];
            const uint32_t nb = 2 * n - 32 * ch < 32u ? 2 * n - 32 * ch : 32u;
            for (uint32_t st = 0; st < nb; st++) bm.push((word >> (31u - st)) & 1u);
        }
        hunt_finish<NWP>(p.out, n, slot0 + lane, bm.c, bm.l);
    }
}

// ---------------------------------------------------------------------------
// front kernel, one warp per candidate, K clocks per pass (n <= 256): the product path.
//
// With rows kept as dense masks the cost of x -> M.x does not depend on how sparse M is, so the warp applies
// P = A^K instead of A: one pass of NW*NW LOP3 + NW votes advances the state K clocks.  The output bits of the
// K-1 clocks in between, u[t+i] = (row 0 of A^i) . x[t] (binary_matrix.rs:72-73 emits x[0] after every step), come
// from one more row group: lane i-1 holds row 0 of A^i as its "extra" row, and one more vote returns all of them.
// A block of K clocks is therefore (NW+1) * (NW + 3) warp instructions, where hunt_front_warp_kernel spends
// K * NW * (NW + 3).  The powers are built by K-1 sparse left multiplications P <- A.P: the rows of P lie in
// shared memory (48-byte row stride: conflict-light 128-bit gathers), lane l XORs the four rows its row of A taps
// (an empty tap slot repeats the previous pick and is skipped), for each of its NW rows.
//
// Persistent warps (one per CTA) pull groups of 32 candidates from an atomic counter.  Generation and
// Berlekamp-Massey run thread-per-candidate around the warp phase as in hunt_front_warp_kernel; the generated
// tap lists wait in a per-warp global scratch ([row][32] words, 24 KB at n = 192, re-used by every group the
// warp takes, so it lives in L2 and shared memory stays small enough for 16 warps per SM).
// shared: pow[n][RS] words, seq[ceil(2n/32)][32] words (bit t of a candidate's sequence = word t/32, bit t%32)
// ---------------------------------------------------------------------------
template <int NW, int K>
__global__ void __launch_bounds__(32, (NW <= 6 ? 16 : 8))
hunt_front_pow_kernel(FrontParams p)
{
    uint32_t *const tap_scratch = p.tap_scratch;
    uint32_t *const group_counter = p.group_counter;
    constexpr int NWP = NW + 1;
    constexpr int RW = NW <= 4 ? 4 : 8;                    // words per row of P in shared memory
    constexpr int RS = NW <= 4 ? 4 : 12;                   // row stride: an odd number of 16-byte units per 4 words read
    static_assert(32 % K == 0 && K >= 2, "K must divide 32");
    extern __shared__ __align__(16) uint32_t hsm[];
    const uint32_t lane = threadIdx.x, n = p.n;
    const uint32_t SW = (2 * n + 31) / 32;                 // sequence words per candidate
    uint32_t *pw = hsm;
    uint32_t *seq = hsm + (size_t)n * RS;
    uint32_t *taps = tap_scratch + (size_t)blockIdx.x * n * 32u;
    const uint32_t ngroups = (p.count + 31u) / 32u;
    const uint64_t keep_pol = l2_keep_policy();

    for (;;) {
        uint32_t g = 0;
        if (lane == 0) g = atomicAdd(group_counter, 1u);
        g = __shfl_sync(0xffffffffu, g, 0);
        if (g >= ngroups) break;
        const uint32_t slot0 = g * 32u;
        const uint32_t ncand = p.count - slot0 < 32u ? p.count - slot0 : 32u;

        // phase 1: lane c generates candidate slot0 + c (binary_matrix.rs:81-101)
        if (lane < ncand) {
            const uint64_t cand = p.first + (uint64_t)(slot0 + lane) * p.stride;
            uint32_t acc = 0, prev = 0;
            hunt_generate_fast<NW>(n, p.seed, cand, [&](uint32_t row, uint32_t k, uint32_t col) {
                const uint32_t cc = col == n ? prev : col;     // empty slot: repeat the previous pick
                prev = cc;
                acc |= cc << (8 * k);
                if (k == 3) { st_keep(taps + row * 32u + lane, acc, keep_pol); acc = 0; }
            });
        }
        __syncwarp();

        // phase 2: one candidate at a time
        for (uint32_t c = 0; c < ncand; c++) {
            uint32_t T[NW], M[NW][NW], E[NW];
#pragma unroll
            for (int w = 0; w < NW; w++) E[w] = 0;
#pragma unroll
            for (int j = 0; j < NW; j++) {
                const uint32_t row = lane + 32u * j;
                T[j] = row < n ? ld_keep(taps + row * 32u + c, keep_pol) : 0xffffffffu;   // rows >= n are never used
#pragma unroll
                for (int w = 0; w < NW; w++) M[j][w] = 0;
                if (row < n) {
#pragma unroll
                    for (int k = 0; k < 4; k++) {
                        const uint32_t t = (T[j] >> (8 * k)) & 0xffu, bit = 1u << (t & 31u);
#pragma unroll
                        for (int w = 0; w < NW; w++) M[j][w] |= (t >> 5) == (uint32_t)w ? bit : 0u;
                    }
                }
            }
            // M = rows of A.  K-1 times: publish P, note row 0 of P = A^i for lane i-1, P <- A.P
#pragma unroll 1
            for (uint32_t i = 1; i < (uint32_t)K; i++) {
#pragma unroll
                for (int j = 0; j < NW; j++) {
                    const uint32_t row = lane + 32u * j;
                    if (row < n) {
                        uint32_t *dst = pw + (size_t)row * RS;
                        if (RW == 4) {
                            *(uint4 *)dst = make_uint4(M[j][0], NW > 1 ? M[j][NW > 1 ? 1 : 0] : 0u, NW > 2 ? M[j][NW > 2 ? 2 : 0] : 0u,
                                                       NW > 3 ? M[j][NW > 3 ? 3 : 0] : 0u);
                        } else {
                            *(uint4 *)dst = make_uint4(M[j][0], M[j][NW > 1 ? 1 : 0], M[j][NW > 2 ? 2 : 0], M[j][NW > 3 ? 3 : 0]);
                            *(uint4 *)(dst + 4) = make_uint4(M[j][NW > 4 ? 4 : 0], NW > 5 ? M[j][NW > 5 ? 5 : 0] : 0u,
                                                             NW > 6 ? M[j][NW > 6 ? 6 : 0] : 0u, NW > 7 ? M[j][NW > 7 ? 7 : 0] : 0u);
                        }
                    }
                }
                __syncwarp();
                {
                    const uint4 e0 = *(const uint4 *)pw;
                    uint4 e1 = make_uint4(0, 0, 0, 0);
                    if (RW == 8) e1 = *(const uint4 *)(pw + 4);
                    const uint32_t ev[8] = {e0.x, e0.y, e0.z, e0.w, e1.x, e1.y, e1.z, e1.w};
                    if (lane == i - 1u) {
#pragma unroll
                        for (int w = 0; w < NW; w++) E[w] = ev[w];
                    }
                }
#pragma unroll
                for (int j = 0; j < NW; j++) {
                    uint32_t a[8] = {0, 0, 0, 0, 0, 0, 0, 0};
                    if (lane + 32u * j < n) {
#pragma unroll
                        for (int k = 0; k < 4; k++) {
                            const uint32_t t = (T[j] >> (8 * k)) & 0xffu;
                            const uint32_t keep = (k == 0 || t != ((T[j] >> (8 * (k > 0 ? k - 1 : 0))) & 0xffu)) ? 0xffffffffu : 0u;
                            const uint32_t *src = pw + (size_t)t * RS;
                            const uint4 v0 = *(const uint4 *)src;
                            a[0] ^= v0.x & keep; a[1] ^= v0.y & keep; a[2] ^= v0.z & keep; a[3] ^= v0.w & keep;
                            if (RW == 8) {
                                const uint4 v1 = *(const uint4 *)(src + 4);
                                a[4] ^= v1.x & keep; a[5] ^= v1.y & keep; a[6] ^= v1.z & keep; a[7] ^= v1.w & keep;
                            }
                        }
                    }
#pragma unroll
                    for (int w = 0; w < NW; w++) M[j][w] = a[w];
                }
                __syncwarp();
            }
            // M = rows of A^K, E (lanes 0..K-2) = row 0 of A^(lane+1)
            uint32_t W[NW];
#pragma unroll
            for (int w = 0; w < NW; w++) {
                const uint32_t lo = (uint32_t)w * 32u;
                W[w] = n >= lo + 32 ? 0xffffffffu : (n > lo ? (1u << (n - lo)) - 1u : 0u);   // b = all ones (rnghunt.rs:28)
            }
            uint32_t mine = 0;
            for (uint32_t ch = 0; ch < SW; ch++) {
                uint32_t acc = 0;
#pragma unroll
                for (int q = 0; q < 32 / K; q++) {
                    uint32_t Wn[NW];
#pragma unroll
                    for (int j = 0; j < NW; j++) {
                        uint32_t v = W[0] & M[j][0];
#pragma unroll
                        for (int w = 1; w < NW; w++) v ^= W[w] & M[j][w];
                        Wn[j] = ballot_odd(__popc(v));
                    }
                    uint32_t x = W[0] & E[0];
#pragma unroll
                    for (int w = 1; w < NW; w++) x ^= W[w] & E[w];
                    const uint32_t X = ballot_odd(__popc(x));                 // bit i-1 = u[t+i], i < K
#pragma unroll
                    for (int w = 0; w < NW; w++) W[w] = Wn[w];
                    acc |= (X | ((W[0] & 1u) << (K - 1))) << (q * K);         // u[t+K] = x[0] of the new state
                }
                if (lane == ch) mine = acc;
            }
            if (lane < SW) seq[lane * 32u + c] = mine;
        }
        __syncwarp();

        // phase 3: lane c runs Berlekamp-Massey over candidate c's 2n bits (berlekamp_massey.rs:5-31)
        if (lane < ncand) {
            HuntBM<NWP> bm;
            bm.init();
            for (uint32_t ch = 0; ch < SW; ch++) {
                const uint32_t word = seq[ch * 32u + lane];
                const uint32_t nb = 2 * n - 32 * ch < 32u ? 2 * n - 32 * ch : 32u;
                for (uint32_t st = 0; st < nb; st++) bm.push((word >> st) & 1u);
            }
            hunt_finish<NWP>(p.out, n, slot0 + lane, bm.c, bm.l);
        }
        __syncwarp();
    }
}

// ---------------------------------------------------------------------------
// front kernel for caller-supplied dense matrices (bbb_rnghunt_eval): one thread per matrix,
// BinaryMatrix layout (binary_matrix.rs:15-19), dot() as the XOR of the selected columns (:53-64)
// ---------------------------------------------------------------------------
template <int NW64>
__global__ void __launch_bounds__(64)
hunt_front_dense_kernel(uint32_t n, uint32_t count, const uint64_t *__restrict__ mats, HuntOut out)
{
    constexpr int NWP = (NW64 * 64 + 1 + 31) / 32;
    const uint32_t slot = blockIdx.x * blockDim.x + threadIdx.x;
    if (slot >= count) return;
    const uint64_t *m = mats + (size_t)slot * n * NW64;
    uint64_t x[NW64];
#pragma unroll
    for (int w = 0; w < NW64; w++) {   // BinaryVector::from_bits(all ones): bit i at word i/64, bit 63 - i%64
        const uint32_t lo = (uint32_t)w * 64u;
        x[w] = n >= lo + 64 ? ~0ull : (n > lo ? ~0ull << (64 - (n - lo)) : 0ull);
    }
    HuntBM<NWP> bm;
    bm.init();
    for (uint32_t step = 0; step < 2 * n; step++) {
        uint64_t y[NW64];
#pragma unroll
        for (int w = 0; w < NW64; w++) y[w] = 0;
#pragma unroll
        for (int w = 0; w < NW64; w++) {
            uint64_t bits = x[w];
            while (bits) {
                const uint32_t lz = __clzll(bits);
                bits &= ~(0x8000000000000000ull >> lz);
                const uint64_t *col = m + (size_t)((uint32_t)w * 64u + lz) * NW64;
#pragma unroll
                for (int v = 0; v < NW64; v++) y[v] ^= col[v];
            }
        }
#pragma unroll
        for (int w = 0; w < NW64; w++) x[w] = y[w];
        bm.push((uint32_t)(x[0] >> 63));
    }
    hunt_finish<NWP>(out, n, slot, bm.c, bm.l);
}

// ---------------------------------------------------------------------------
// primitivity: x^e mod p == 1 ?  (binary_polynomial.rs:136-175, check_integer)
// polynomials LSB-first in NWP 32-bit words, deg p = n
// ---------------------------------------------------------------------------
template <int NWP>
__device__ __forceinline__ void poly_mulx(uint32_t (&v)[NWP], const uint32_t (&p)[NWP], uint32_t n)
{
#pragma unroll
    for (int w = NWP - 1; w > 0; w--) v[w] = (v[w] << 1) | (v[w - 1] >> 31);
    v[0] <<= 1;
    // bit n set -> subtract p (clears it)
    uint32_t top = 0;
#pragma unroll
    for (int w = 0; w < NWP; w++) if ((uint32_t)w == (n >> 5)) top = (v[w] >> (n & 31)) & 1u;
    const uint32_t m = 0u - top;
#pragma unroll
    for (int w = 0; w < NWP; w++) v[w] ^= p[w] & m;
}

// r = r * r mod p, Horner over the bits of r from the top (binary_polynomial.rs:88-133 computes the
// same product by shift-and-add)
template <int NWP>
__device__ __forceinline__ void poly_square(uint32_t (&r)[NWP], const uint32_t (&p)[NWP], uint32_t n)
{
    uint32_t acc[NWP];
#pragma unroll
    for (int w = 0; w < NWP; w++) acc[w] = 0;
#pragma unroll
    for (int w = NWP - 1; w >= 0; w--) {
        const uint32_t base = (uint32_t)w * 32u;
        if (base < n) {                                   // r has degree < n
            const uint32_t word = r[w];
            const int hi = n - 1 - base < 31u ? (int)(n - 1 - base) : 31;
#pragma unroll 1
            for (int b = hi; b >= 0; b--) {
                poly_mulx<NWP>(acc, p, n);
                const uint32_t m = 0u - ((word >> b) & 1u);
#pragma unroll
                for (int v = 0; v < NWP; v++) acc[v] ^= r[v] & m;
            }
        }
    }
#pragma unroll
    for (int w = 0; w < NWP; w++) r[w] = acc[w];
}

struct CheckParams {
    uint32_t n, nexp;                  // nexp exponents per polynomial in this launch
    uint32_t exp_first;                // index of the first exponent (0 = 2^n-1, i >= 1 = cofactor i)
    const uint32_t *npoly;             // device: how many polynomials (survivors, or order passers)
    const uint32_t *poly_index;        // [npoly] survivor index of polynomial q, or null = identity
    const uint32_t *polys;             // [..][NWP]
    const uint64_t *exp_words;         // all exponents of n back to back, MS word first
    const uint32_t *exp_off;           // [nf+1] word offsets into exp_words
    uint32_t *result;                  // [survivor] bit (exp index) set when x^e == 1
};

// The item counts live on the device (the front kernel's survivor counter, the compaction's pass counter), so
// the host never waits for them: every check launch is a fixed grid striding over *npoly * nexp items.
template <int NWP>
__global__ void __launch_bounds__(128)
hunt_check_kernel(CheckParams q)
{
    const uint32_t nitems = *q.npoly * q.nexp;
    for (uint32_t item = blockIdx.x * blockDim.x + threadIdx.x; item < nitems; item += gridDim.x * blockDim.x) {
        const uint32_t pi = item / q.nexp, ei = q.exp_first + item % q.nexp;
        const uint32_t sidx = q.poly_index ? q.poly_index[pi] : pi;
        uint32_t p[NWP], r[NWP];
#pragma unroll
        for (int w = 0; w < NWP; w++) { p[w] = q.polys[(size_t)sidx * NWP + w]; r[w] = 0; }
        r[0] = 1;                                              // modexp starts from 1 (:138-142)
        const uint64_t *e = q.exp_words + q.exp_off[ei];
        const uint32_t ew = q.exp_off[ei + 1] - q.exp_off[ei];
        bool started = false;
        for (uint32_t w = 0; w < ew; w++) {
            const uint64_t word = e[w];
            for (int b = 63; b >= 0; b--) {
                const uint32_t bit = (uint32_t)(word >> b) & 1u;
                if (!started) {
                    if (!bit) continue;                        // 1 squared stays 1
                    started = true;
                } else {
                    poly_square<NWP>(r, p, q.n);
                }
                if (bit) poly_mulx<NWP>(r, p, q.n);
            }
        }
        uint32_t rest = r[0] ^ 1u;
#pragma unroll
        for (int w = 1; w < NWP; w++) rest |= r[w];
        if (rest == 0) atomicOr(&q.result[sidx], 1u << (ei > 31 ? 31 : ei));
    }
}

// The same test for n = 32*NW with a per-polynomial table R_j = x^(n+j) mod p, j < 32, in shared
// memory (word [j][w][thread]): squaring = bit spread + folding the high half 32 bits at a time
// (one table row per set bit), about 4x fewer instructions than shift-and-add.
__device__ __forceinline__ uint32_t spread16(uint32_t x)   // bit i of the low half -> bit 2i
{
    x &= 0xffffu;
    x = (x | (x << 8)) & 0x00ff00ffu;
    x = (x | (x << 4)) & 0x0f0f0f0fu;
    x = (x | (x << 2)) & 0x33333333u;
    x = (x | (x << 1)) & 0x55555555u;
    return x;
}

template <int NW>
__global__ void __launch_bounds__(64)
hunt_check32_kernel(CheckParams q, uint32_t stride)
{
    extern __shared__ uint32_t rsm[];
    const uint32_t B = blockDim.x, tid = threadIdx.x;
    const uint32_t nitems = *q.npoly * q.nexp;
    uint32_t *const R = rsm + tid;
    for (uint32_t item = blockIdx.x * B + tid; item < nitems; item += gridDim.x * B) {
        const uint32_t pi = item / q.nexp, ei = q.exp_first + item % q.nexp;
        const uint32_t sidx = q.poly_index ? q.poly_index[pi] : pi;
        uint32_t pl[NW], r[NW];                        // p without its leading term = x^n mod p
#pragma unroll
        for (int w = 0; w < NW; w++) pl[w] = q.polys[(size_t)sidx * stride + w];
        {
            uint32_t t[NW];
#pragma unroll
            for (int w = 0; w < NW; w++) t[w] = pl[w];
            for (uint32_t j = 0; j < 32; j++) {
#pragma unroll
                for (int w = 0; w < NW; w++) R[(j * NW + w) * B] = t[w];
                const uint32_t m = 0u - (t[NW - 1] >> 31);
#pragma unroll
                for (int w = NW - 1; w > 0; w--) t[w] = ((t[w] << 1) | (t[w - 1] >> 31)) ^ (pl[w] & m);
                t[0] = (t[0] << 1) ^ (pl[0] & m);
            }
        }
#pragma unroll
        for (int w = 0; w < NW; w++) r[w] = 0;
        r[0] = 1;
        const uint64_t *e = q.exp_words + q.exp_off[ei];
        const uint32_t ew = q.exp_off[ei + 1] - q.exp_off[ei];
        bool started = false;
        for (uint32_t wi = 0; wi < ew; wi++) {
            const uint64_t word = e[wi];
            for (int b = 63; b >= 0; b--) {
                const uint32_t bit = (uint32_t)(word >> b) & 1u;
                if (!started) {
                    if (!bit) continue;
                    started = true;
                } else {
                    uint32_t V[2 * NW];
#pragma unroll
                    for (int w = 0; w < NW; w++) { V[2 * w] = spread16(r[w]); V[2 * w + 1] = spread16(r[w] >> 16); }
#pragma unroll
                    for (int qq = NW - 1; qq >= 0; qq--) {
                        uint32_t chunk = V[NW + qq];               // coefficients of x^(n+32qq+j)
                        while (chunk) {
                            const uint32_t j = __ffs(chunk) - 1;
                            chunk &= chunk - 1;
                            const uint32_t *Rj = R + (size_t)j * NW * B;
#pragma unroll
                            for (int w = 0; w < NW; w++) V[qq + w] ^= Rj[w * B];
                        }
                    }
#pragma unroll
                    for (int w = 0; w < NW; w++) r[w] = V[w];
                }
                if (bit) {                                         // times x
                    const uint32_t m = 0u - (r[NW - 1] >> 31);
#pragma unroll
                    for (int w = NW - 1; w > 0; w--) r[w] = ((r[w] << 1) | (r[w - 1] >> 31)) ^ (pl[w] & m);
                    r[0] = (r[0] << 1) ^ (pl[0] & m);
                }
            }
        }
        uint32_t rest = r[0] ^ 1u;
#pragma unroll
        for (int w = 1; w < NW; w++) rest |= r[w];
        if (rest == 0) atomicOr(&q.result[sidx], 1u << (ei > 31 ? 31 : ei));
    }
}

// survivors whose x^(2^n-1) == 1 go on to the cofactor tests (binary_polynomial.rs:198-208)
__global__ void hunt_compact_kernel(const uint32_t *__restrict__ result, const uint32_t *nsurv_p, uint32_t *pass_count,
                                    uint32_t *pass_index)
{
    const uint32_t nsurv = *nsurv_p;
    for (uint32_t s = blockIdx.x * blockDim.x + threadIdx.x; s < nsurv; s += gridDim.x * blockDim.x)
        if (result[s] & 1u) pass_index[atomicAdd(pass_count, 1u)] = s;
}

// Final stage of every survivor; accepted candidates go to the hit list, which lives in pinned HOST memory
// (zero-copy: hits are rare, 1 in 2240 at n = 192) together with their polynomials when the caller wants them.
__global__ void hunt_classify_kernel(const uint32_t *__restrict__ result, const uint32_t *__restrict__ surv_cand,
                                     const uint32_t *nsurv_p, uint8_t *stage, uint32_t *hit_count, uint32_t hit_cap,
                                     uint32_t *hits, const uint32_t *__restrict__ surv_poly, uint32_t nwp, uint32_t *hit_poly)
{
    const uint32_t nsurv = *nsurv_p;
    for (uint32_t s = blockIdx.x * blockDim.x + threadIdx.x; s < nsurv; s += gridDim.x * blockDim.x) {
        const uint32_t r = result[s];
        uint8_t st;
        if (!(r & 1u)) st = BBB_STAGE_ORDER;
        else if (r & ~1u) st = BBB_STAGE_FACTOR;
        else st = BBB_STAGE_PRIMITIVE;
        const uint32_t slot = surv_cand[s];
        stage[slot] = st;
        if (st == BBB_STAGE_PRIMITIVE) {
            const uint32_t h = atomicAdd(hit_count, 1u);
            if (h < hit_cap) {
                hits[h] = slot;
                if (hit_poly)
                    for (uint32_t w = 0; w < nwp; w++) hit_poly[(size_t)h * nwp + w] = surv_poly[(size_t)s * nwp + w];
            }
        }
    }
}

__global__ void hunt_count_kernel(const uint8_t *__restrict__ stage, uint32_t count, unsigned long long *counts)
{
    __shared__ uint32_t sh[BBB_NSTAGES];
    if (threadIdx.x < BBB_NSTAGES) sh[threadIdx.x] = 0;
    __syncthreads();
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < count; i += gridDim.x * blockDim.x) {
        const uint8_t s = stage[i];
        if (s < BBB_NSTAGES) atomicAdd(&sh[s], 1u);
    }
    __syncthreads();
    if (threadIdx.x < BBB_NSTAGES && sh[threadIdx.x]) atomicAdd(&counts[threadIdx.x], (unsigned long long)sh[threadIdx.x]);
}

// ---------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------
// One batch in flight.  Two of them alternate between the ctx's two streams, so the short, badly filled
// back-end kernels of batch i run beside the front kernel of batch i+1, and the host reads batch i's hit
// list (pinned memory, written by the classify kernel) only after it has queued batch i+1.
struct HuntSlot {
    cudaStream_t stream;
    cudaEvent_t done;
    uint8_t *stage;
    uint32_t *counters;            // device: [0] survivors, [1] order passers, [2] hits
    uint32_t *surv_cand, *surv_poly, *result, *pass_index;
    uint32_t *h_hdr;               // pinned host: copy of counters[0..3]
    uint32_t *h_hits;              // pinned host, device-visible: [hit_cap] accepted slots
    uint32_t *h_hit_poly;          // pinned host: [hit_cap][nwp], or null
    uint32_t hit_cap;
    uint64_t first;                // candidate index of slot 0
    uint32_t nb;                   // candidates in the batch in flight (0 = idle)
};

struct HuntShared {
    unsigned long long *counts;    // [BBB_NSTAGES], accumulated by both streams
    uint32_t *exp_off;
    uint64_t *exp_words;
    uint32_t nf;                   // exponents for n (entry 0 = 2^n - 1)
};

static size_t align256(size_t v) { return (v + 255) & ~(size_t)255; }

// Carves scratch slot 6 (device) and pinned slot 0 (host) for `nslots` batches of `cap` candidates and uploads
// the exponent table of n (factors.rs:7-16: [2^n-1, (2^n-1)/q_1, ...], MS word first).
static int hunt_setup(bbb_ctx *c, uint32_t n, uint32_t cap, uint32_t nwp, bool want_polys, int nslots, HuntShared &sh,
                      HuntSlot *slots)
{
    const uint32_t e0 = BBB_MF_OFFSETS[n - 1], e1 = BBB_MF_OFFSETS[n];
    sh.nf = e1 - e0;
    std::vector<uint32_t> off(sh.nf + 1);
    std::vector<uint64_t> words;
    for (uint32_t i = 0; i < sh.nf; i++) {
        off[i] = (uint32_t)words.size();
        for (uint32_t w = BBB_MF_ENTRIES[e0 + i]; w < BBB_MF_ENTRIES[e0 + i + 1]; w++) words.push_back(BBB_MF_WORDS[w]);
    }
    off[sh.nf] = (uint32_t)words.size();
    const uint32_t hit_cap = cap;                  // sized for the worst case: every candidate of a batch accepted
    size_t at = 0, hat = 0;
    auto take = [&](size_t bytes) { size_t o = at; at += align256(bytes); return o; };
    auto htake = [&](size_t bytes) { size_t o = hat; hat += align256(bytes); return o; };
    const size_t o_counts = take(8 * BBB_NSTAGES), o_eoff = take(off.size() * 4), o_ew = take(words.size() * 8);
    size_t o_stage[2], o_ctr[2], o_cand[2], o_poly[2], o_res[2], o_pass[2], h_hdr[2], h_hits[2], h_poly[2];
    for (int i = 0; i < nslots; i++) {
        o_stage[i] = take(cap); o_ctr[i] = take(16); o_cand[i] = take((size_t)cap * 4);
        o_poly[i] = take((size_t)cap * nwp * 4); o_res[i] = take((size_t)cap * 4); o_pass[i] = take((size_t)cap * 4);
        h_hdr[i] = htake(16); h_hits[i] = htake((size_t)hit_cap * 4);
        h_poly[i] = want_polys ? htake((size_t)hit_cap * nwp * 4) : 0;
    }
    char *d = (char *)c->scratch(6, at);
    char *h = (char *)c->pinned(0, hat);
    if (!d || !h) return BBB_ECUDA;
    sh.counts = (unsigned long long *)(d + o_counts);
    sh.exp_off = (uint32_t *)(d + o_eoff);
    sh.exp_words = (uint64_t *)(d + o_ew);
    for (int i = 0; i < nslots; i++) {
        HuntSlot &b = slots[i];
        b.stream = i ? c->stream2 : c->stream;
        b.done = c->ev_slot[i];
        b.stage = (uint8_t *)(d + o_stage[i]);
        b.counters = (uint32_t *)(d + o_ctr[i]);
        b.surv_cand = (uint32_t *)(d + o_cand[i]);
        b.surv_poly = (uint32_t *)(d + o_poly[i]);
        b.result = (uint32_t *)(d + o_res[i]);
        b.pass_index = (uint32_t *)(d + o_pass[i]);
        b.h_hdr = (uint32_t *)(h + h_hdr[i]);
        b.h_hits = (uint32_t *)(h + h_hits[i]);
        b.h_hit_poly = want_polys ? (uint32_t *)(h + h_poly[i]) : nullptr;
        b.hit_cap = hit_cap;
        b.first = 0;
        b.nb = 0;
    }
    BBB_CUDA(c, c->h2d(sh.exp_off, off.data(), off.size() * 4));
    BBB_CUDA(c, c->h2d(sh.exp_words, words.data(), words.size() * 8));
    BBB_CUDA(c, cudaMemsetAsync(sh.counts, 0, 8 * BBB_NSTAGES, c->stream));
    BBB_CUDA(c, cudaStreamSynchronize(c->stream));   // off / words are about to go out of scope
    return BBB_OK;
}

static int launch_check(bbb_ctx *c, cudaStream_t st, uint32_t nwp, const CheckParams &q, uint64_t max_items)
{
    if (!max_items) return BBB_OK;
    if (q.n % 32 == 0 && !c->tune_hunt_slow_check) {
        const uint32_t nw = q.n / 32, B = 64;
        const size_t smem = (size_t)32 * nw * B * 4;
        const uint64_t want = (max_items + B - 1) / B, most = (uint64_t)c->sm_count * (smem > 56 * 1024 ? 2 : 4);
        const uint32_t blocks = (uint32_t)(want < most ? want : most);
#define CHECK32_CASE(N)                                                                                        \
    case N:                                                                                                    \
        BBB_CUDA(c, cudaFuncSetAttribute(hunt_check32_kernel<N>, cudaFuncAttributeMaxDynamicSharedMemorySize,   \
                                         (int)smem));                                                          \
        hunt_check32_kernel<N><<<blocks, B, smem, st>>>(q, nwp);                                               \
        break
        switch (nw) {
            CHECK32_CASE(1); CHECK32_CASE(2); CHECK32_CASE(3); CHECK32_CASE(4); CHECK32_CASE(5); CHECK32_CASE(6);
            CHECK32_CASE(7); CHECK32_CASE(8); CHECK32_CASE(9); CHECK32_CASE(10); CHECK32_CASE(11); CHECK32_CASE(12);
            CHECK32_CASE(13); CHECK32_CASE(14); CHECK32_CASE(15); CHECK32_CASE(16);
        default: return c->fail(BBB_EINVAL, "bad n");
        }
#undef CHECK32_CASE
        c->last_launches++;
        BBB_CUDA(c, cudaGetLastError());
        return BBB_OK;
    }
    const uint64_t want = (max_items + 127) / 128, most = (uint64_t)c->sm_count * 8;
    const uint32_t blocks = (uint32_t)(want < most ? want : most);
#define CHECK_CASE(N) case N: hunt_check_kernel<N><<<blocks, 128, 0, st>>>(q); break
    switch (nwp) {
        CHECK_CASE(2); CHECK_CASE(3); CHECK_CASE(4); CHECK_CASE(5); CHECK_CASE(6); CHECK_CASE(7); CHECK_CASE(8);
        CHECK_CASE(9); CHECK_CASE(10); CHECK_CASE(11); CHECK_CASE(12); CHECK_CASE(13); CHECK_CASE(14);
        CHECK_CASE(15); CHECK_CASE(16); CHECK_CASE(17);
    default: return c->fail(BBB_EINVAL, "bad polynomial width");
    }
#undef CHECK_CASE
    c->last_launches++;
    BBB_CUDA(c, cudaGetLastError());
    return BBB_OK;
}

// Stages 2 and 3 of the pipeline for the survivors the front kernel left in `b`, queued on b.stream without a
// single host round trip: the survivor / passer counts are read by the kernels themselves.  Afterwards b.stage
// holds the final verdict of every candidate of the batch, the pinned hit list the accepted slots.
static int hunt_back_end(bbb_ctx *c, uint32_t n, uint32_t nb, uint32_t nwp, const HuntShared &sh, HuntSlot &b)
{
    cudaStream_t st = b.stream;
    BBB_CUDA(c, cudaMemsetAsync(b.result, 0, (size_t)nb * 4, st));
    CheckParams q;
    q.n = n; q.polys = b.surv_poly; q.exp_words = sh.exp_words; q.exp_off = sh.exp_off; q.result = b.result;
    q.npoly = b.counters; q.nexp = 1; q.exp_first = 0; q.poly_index = nullptr;            // x^(2^n-1) == 1 ?
    int rc = launch_check(c, st, nwp, q, nb);
    if (rc) return rc;
    const uint32_t gb = std::min<uint32_t>((nb + 255) / 256, (uint32_t)c->sm_count * 4);
    if (sh.nf > 1) {
        hunt_compact_kernel<<<gb, 256, 0, st>>>(b.result, b.counters, b.counters + 1, b.pass_index);
        c->last_launches++;
        q.npoly = b.counters + 1; q.nexp = sh.nf - 1; q.exp_first = 1; q.poly_index = b.pass_index;
        rc = launch_check(c, st, nwp, q, (uint64_t)nb * (sh.nf - 1));                     // cofactors
        if (rc) return rc;
    }
    hunt_classify_kernel<<<gb, 256, 0, st>>>(b.result, b.surv_cand, b.counters, b.stage, b.counters + 2, b.hit_cap,
                                            b.h_hits, b.surv_poly, nwp, b.h_hit_poly);
    hunt_count_kernel<<<c->sm_count * 2, 256, 0, st>>>(b.stage, nb, sh.counts);
    c->last_launches += 2;
    BBB_CUDA(c, cudaGetLastError());
    BBB_CUDA(c, cudaMemcpyAsync(b.h_hdr, b.counters, 16, cudaMemcpyDeviceToHost, st));
    c->d2h_bytes += 16;
    BBB_CUDA(c, cudaEventRecord(b.done, st));
    b.nb = nb;
    return BBB_OK;
}

struct HuntHit {
    uint64_t cand;
    std::vector<uint32_t> poly;    // nwp words, bit i = coefficient of x^i (empty unless requested)
};

// Waits for the batch in `b` and appends its accepted candidates.
static int hunt_collect(bbb_ctx *c, HuntSlot &b, uint64_t stride, uint32_t nwp, std::vector<HuntHit> &found)
{
    if (!b.nb) return BBB_OK;
    BBB_CUDA(c, cudaEventSynchronize(b.done));
    const uint32_t nh = b.h_hdr[2];
    b.nb = 0;
    if (nh > b.hit_cap) return c->fail(BBB_EINVAL, "%u hits in one batch: lower hunt_batch", nh);
    c->d2h_bytes += (uint64_t)nh * 4 * (b.h_hit_poly ? 1 + nwp : 1);
    for (uint32_t i = 0; i < nh; i++) {
        HuntHit h;
        h.cand = b.first + (uint64_t)b.h_hits[i] * stride;
        if (b.h_hit_poly) h.poly.assign(b.h_hit_poly + (size_t)i * nwp, b.h_hit_poly + (size_t)(i + 1) * nwp);
        found.push_back(std::move(h));
    }
    return BBB_OK;
}

// persistent warps of hunt_front_pow_kernel: one per CTA, up to 16 per SM
static uint32_t hunt_pow_grid(const bbb_ctx *c, uint32_t count)
{
    const uint32_t groups = (count + 31u) / 32u, most = (uint32_t)c->sm_count * 16u;   // n > 192: 8 are resident, the rest queue
    return groups < most ? (groups ? groups : 1u) : most;
}

static int launch_front(bbb_ctx *c, cudaStream_t st, const FrontParams &p)
{
    const uint32_t n = p.n, nw = (n + 31) / 32;
    // n <= 256: warp per candidate, K clocks per pass ("hunt_pow_k" = 4 or 8); "hunt_smem_taps" = 2 selects the
    // one-clock-per-pass warp kernel, 1 the thread-per-candidate kernel (tests: independent implementations)
    if (!p.dump_taps && nw <= 8 && c->tune_hunt_smem_taps == 0 && p.tap_scratch) {
        const uint32_t K = c->tune_hunt_pow_k == 4 ? 4 : 8;
        const uint32_t rs = nw <= 4 ? 4 : 12;
        const size_t smem = ((size_t)n * rs + (size_t)((2 * n + 31) / 32) * 32) * 4;
        const uint32_t blocks = hunt_pow_grid(c, p.count);
#define FRONT_POW_CASE(NW)                                                                                          \
    case NW:                                                                                                        \
        if (K == 8) {                                                                                               \
            BBB_CUDA(c, cudaFuncSetAttribute(hunt_front_pow_kernel<NW, 8>, cudaFuncAttributeMaxDynamicSharedMemorySize, \
                                             (int)smem));                                                           \
            hunt_front_pow_kernel<NW, 8><<<blocks, 32, smem, st>>>(p);                                              \
        } else {                                                                                                    \
            BBB_CUDA(c, cudaFuncSetAttribute(hunt_front_pow_kernel<NW, 4>, cudaFuncAttributeMaxDynamicSharedMemorySize, \
                                             (int)smem));                                                           \
            hunt_front_pow_kernel<NW, 4><<<blocks, 32, smem, st>>>(p);                                              \
        }                                                                                                           \
        break
        switch (nw) {
            FRONT_POW_CASE(1); FRONT_POW_CASE(2); FRONT_POW_CASE(3); FRONT_POW_CASE(4); FRONT_POW_CASE(5);
            FRONT_POW_CASE(6); FRONT_POW_CASE(7); FRONT_POW_CASE(8);
        default: break;
        }
#undef FRONT_POW_CASE
        c->last_launches++;
        BBB_CUDA(c, cudaGetLastError());
        return BBB_OK;
    }
    if (!p.dump_taps && nw <= 8 && c->tune_hunt_smem_taps != 1) {
        const uint32_t cpw = c->tune_hunt_cpw == 16 ? 16 : 32;
        const size_t smem = ((size_t)n * (cpw + 1) + (size_t)((2 * n + 31) / 32) * cpw) * 4;
        const uint32_t blocks = (p.count + cpw - 1) / cpw;
#define FRONT_WARP_CASE(NW)                                                                                         \
    case NW:                                                                                                        \
        if (cpw == 32) {                                                                                            \
            BBB_CUDA(c, cudaFuncSetAttribute(hunt_front_warp_kernel<NW, 32>, cudaFuncAttributeMaxDynamicSharedMemorySize, \
                                             (int)smem));                                                           \
            hunt_front_warp_kernel<NW, 32><<<blocks, 32, smem, st>>>(p);                                            \
        } else {                                                                                                    \
            BBB_CUDA(c, cudaFuncSetAttribute(hunt_front_warp_kernel<NW, 16>, cudaFuncAttributeMaxDynamicSharedMemorySize, \
                                             (int)smem));                                                           \
            hunt_front_warp_kernel<NW, 16><<<blocks, 32, smem, st>>>(p);                                            \
        }                                                                                                           \
        break
        switch (nw) {
            FRONT_WARP_CASE(1); FRONT_WARP_CASE(2); FRONT_WARP_CASE(3); FRONT_WARP_CASE(4); FRONT_WARP_CASE(5);
            FRONT_WARP_CASE(6); FRONT_WARP_CASE(7); FRONT_WARP_CASE(8);
        default: break;
        }
#undef FRONT_WARP_CASE
        c->last_launches++;
        BBB_CUDA(c, cudaGetLastError());
        return BBB_OK;
    }
    const uint32_t B = 32;
    const bool wide = n > 255;                         // the "no tap" index n must fit the tap type
    const size_t smem = (size_t)B * ((n + 1) * 4 + (size_t)n * 4 * (wide ? 2 : 1));
    const uint32_t blocks = (p.count + B - 1) / B;
#define FRONT_CASE(NW, T)                                                                              \
    case NW:                                                                                           \
        BBB_CUDA(c, cudaFuncSetAttribute(hunt_front_kernel<NW, T>, cudaFuncAttributeMaxDynamicSharedMemorySize, \
                                         (int)smem));                                                  \
        hunt_front_kernel<NW, T><<<blocks, B, smem, st>>>(p);                                          \
        break
    if (!wide) {
        switch (nw) {
            FRONT_CASE(1, uint8_t); FRONT_CASE(2, uint8_t); FRONT_CASE(3, uint8_t); FRONT_CASE(4, uint8_t);
            FRONT_CASE(5, uint8_t); FRONT_CASE(6, uint8_t); FRONT_CASE(7, uint8_t); FRONT_CASE(8, uint8_t);
        default: return c->fail(BBB_EINVAL, "bad n");
        }
    } else {
        switch (nw) {
            FRONT_CASE(8, uint16_t); FRONT_CASE(9, uint16_t); FRONT_CASE(10, uint16_t); FRONT_CASE(11, uint16_t);
            FRONT_CASE(12, uint16_t); FRONT_CASE(13, uint16_t); FRONT_CASE(14, uint16_t); FRONT_CASE(15, uint16_t);
            FRONT_CASE(16, uint16_t);
        default: return c->fail(BBB_EINVAL, "bad n");
        }
    }
#undef FRONT_CASE
    c->last_launches++;
    BBB_CUDA(c, cudaGetLastError());
    return BBB_OK;
}

static int check_hunt_n(bbb_ctx *c, uint32_t n)
{
    if (!c) return BBB_EINVAL;
    if (n < 8 || n > BBB_MAX_N) return c->fail(BBB_EINVAL, "n=%u: need 8 <= n <= 512", n);
    return BBB_OK;
}

// p(x) as nwp LSB-first words -> BinaryPolynomial words: coefficient j (of x^(n-j)) at word j/64, bit 63 - j%64
static void hunt_poly_to_reference(uint32_t n, const uint32_t *p, uint64_t *out, uint32_t *weight)
{
    const uint32_t pw64 = (n + 1 + 63) / 64;
    uint32_t ones = 0;
    for (uint32_t w = 0; w < pw64; w++) out[w] = 0;
    for (uint32_t j = 0; j <= n; j++) {
        const uint32_t i = n - j;
        if ((p[i >> 5] >> (i & 31)) & 1u) { out[j >> 6] |= 1ull << (63 - (j & 63)); ones++; }
    }
    if (weight) *weight = ones;
}

static int hunt_sweep(bbb_ctx *c, uint32_t n, uint64_t seed, uint64_t first, uint64_t stride, uint64_t count,
                      uint32_t k, uint64_t *best, uint32_t *nfound, uint64_t *stage_counts, uint8_t *stage_out,
                      uint64_t *poly_words, uint32_t *weights)
{
    int rc = check_hunt_n(c, n);
    if (rc) return rc;
    if (!stride) return c->fail(BBB_EINVAL, "stride must be >= 1");
    if (k && !best) return c->fail(BBB_EINVAL, "best is null");
    BBB_CUDA(c, cudaSetDevice(c->device));
    c->begin_call();
    c->last_kernel = "rnghunt_sweep";
    const uint32_t nwp = (n + 31) / 32 + 1;
    const bool want_polys = poly_words || weights;
    // batches of 2^20 candidates, but at least four of them per call where the call is long enough (a rank's share of a
    // strong-scaled sweep): the back end of a batch hides behind the front kernel of the next one
    uint64_t batch_max = c->tune_hunt_batch ? c->tune_hunt_batch : (1u << 20);
    if (!c->tune_hunt_batch && count < 4 * batch_max) {
        batch_max = ((count + 3) / 4 + 4095) & ~(uint64_t)4095;
        if (batch_max < (1u << 16)) batch_max = 1u << 16;
    }
    const uint32_t cap = (uint32_t)(count < batch_max ? (count ? count : 1) : batch_max);
    HuntShared sh;
    HuntSlot slots[2];
    rc = hunt_setup(c, n, cap, nwp, want_polys, 2, sh, slots);
    if (rc) return rc;
    // the second stream joins the timed region of the first
    BBB_CUDA(c, cudaEventRecord(c->ev_fork, c->stream));
    BBB_CUDA(c, cudaStreamWaitEvent(c->stream2, c->ev_fork, 0));
    std::vector<HuntHit> found;
    uint32_t turn = 0;
    for (uint64_t done = 0; done < count; done += cap, turn ^= 1u) {
        HuntSlot &b = slots[turn];
        rc = hunt_collect(c, b, stride, nwp, found);          // the batch that used this slot two turns ago
        if (rc) return rc;
        const uint32_t nb = (uint32_t)(count - done < cap ? count - done : cap);
        BBB_CUDA(c, cudaMemsetAsync(b.counters, 0, 16, b.stream));
        FrontParams p;
        p.n = n; p.seed = seed; p.first = first + done * stride; p.stride = stride; p.count = nb;
        p.out.stage = b.stage; p.out.surv_count = b.counters; p.out.surv_cand = b.surv_cand;
        p.out.surv_poly = b.surv_poly; p.out.poly_out = nullptr; p.dump_taps = nullptr;
        p.group_counter = b.counters + 3;
        p.tap_scratch = nullptr;
        if (n <= 256) {      // per-slot scratch: the two slots' front kernels may overlap
            const size_t per_slot = (size_t)hunt_pow_grid(c, cap) * n * 32 * 4;
            char *ts = (char *)c->scratch(8, 2 * per_slot);
            if (!ts) return BBB_ECUDA;
            p.tap_scratch = (uint32_t *)(ts + turn * per_slot);
        }
        b.first = p.first;
        rc = launch_front(c, b.stream, p);
        if (rc) return rc;
        rc = hunt_back_end(c, n, nb, nwp, sh, b);
        if (rc) return rc;
        if (stage_out) {
            BBB_CUDA(c, cudaMemcpyAsync(stage_out + done, b.stage, nb, cudaMemcpyDeviceToHost, b.stream));
            c->d2h_bytes += nb;
            BBB_CUDA(c, cudaStreamSynchronize(b.stream));
        }
    }
    for (uint32_t i = 0; i < 2; i++, turn ^= 1u) {
        rc = hunt_collect(c, slots[turn], stride, nwp, found);
        if (rc) return rc;
    }
    BBB_CUDA(c, cudaEventRecord(c->ev_join, c->stream2));
    BBB_CUDA(c, cudaStreamWaitEvent(c->stream, c->ev_join, 0));
    c->end_call();
    unsigned long long counts[BBB_NSTAGES];
    BBB_CUDA(c, c->d2h(counts, sh.counts, sizeof counts));
    BBB_CUDA(c, cudaStreamSynchronize(c->stream));
    if (stage_counts) for (int i = 0; i < BBB_NSTAGES; i++) stage_counts[i] += counts[i];
    std::sort(found.begin(), found.end(), [](const HuntHit &a, const HuntHit &b) { return a.cand < b.cand; });
    const uint32_t nf = (uint32_t)(found.size() < k ? found.size() : k);
    const uint32_t pw64 = (n + 1 + 63) / 64;
    for (uint32_t i = 0; i < nf; i++) {
        best[i] = found[i].cand;
        if (want_polys) {
            uint64_t tmp[(BBB_MAX_N + 64) / 64 + 1];
            uint32_t wt = 0;
            hunt_poly_to_reference(n, found[i].poly.data(), tmp, &wt);
            if (poly_words) memcpy(poly_words + (size_t)i * pw64, tmp, pw64 * 8);
            if (weights) weights[i] = wt;
        }
    }
    if (nfound) *nfound = nf;
    return BBB_OK;
}

extern "C" {

int bbb_rnghunt_sweep(bbb_ctx *c, uint32_t n, uint64_t seed, uint64_t first, uint64_t stride, uint64_t count,
                      uint32_t k, uint64_t *best, uint32_t *nfound, uint64_t *stage_counts, uint8_t *stage_out)
{
    return hunt_sweep(c, n, seed, first, stride, count, k, best, nfound, stage_counts, stage_out, nullptr, nullptr);
}

int bbb_rnghunt_sweep_polys(bbb_ctx *c, uint32_t n, uint64_t seed, uint64_t first, uint64_t stride, uint64_t count,
                            uint32_t k, uint64_t *best, uint32_t *nfound, uint64_t *stage_counts,
                            uint64_t *poly_words, uint32_t *weights)
{
    if (c && !poly_words && !weights) return c->fail(BBB_EINVAL, "poly_words and weights are both null");
    return hunt_sweep(c, n, seed, first, stride, count, k, best, nfound, stage_counts, nullptr, poly_words, weights);
}

int bbb_rnghunt_candidate(bbb_ctx *c, uint32_t n, uint64_t seed, uint64_t cand, uint64_t *matrix_words)
{
    int rc = check_hunt_n(c, n);
    if (rc) return rc;
    if (!matrix_words) return c->fail(BBB_EINVAL, "matrix_words is null");
    BBB_CUDA(c, cudaSetDevice(c->device));
    c->begin_call();
    c->last_kernel = "rnghunt_candidate";
    uint16_t *d_taps = (uint16_t *)c->scratch(5, (size_t)n * 4 * 2);
    if (!d_taps) return BBB_ECUDA;
    FrontParams p;
    memset(&p, 0, sizeof p);
    p.n = n; p.seed = seed; p.first = cand; p.stride = 1; p.count = 1; p.dump_taps = d_taps;
    rc = launch_front(c, c->stream, p);
    if (rc) return rc;
    c->end_call();
    std::vector<uint16_t> taps((size_t)n * 4);
    BBB_CUDA(c, c->d2h(taps.data(), d_taps, taps.size() * 2));
    BBB_CUDA(c, cudaStreamSynchronize(c->stream));
    const uint32_t wpc = (n + 63) / 64;
    memset(matrix_words, 0, (size_t)n * wpc * 8);
    for (uint32_t r = 0; r < n; r++)
        for (uint32_t s = 0; s < 4; s++) {
            const uint16_t col = taps[r * 4 + s];
            if (col == 0xFFFFu) continue;
            matrix_words[(size_t)col * wpc + r / 64] |= 1ull << (63 - r % 64);   // binary_matrix.rs:15-19
        }
    return BBB_OK;
}

int bbb_rnghunt_eval(bbb_ctx *c, uint32_t n, uint64_t ncand, const uint64_t *matrices, uint8_t *stage,
                     uint64_t *poly_words)
{
    int rc = check_hunt_n(c, n);
    if (rc) return rc;
    if (!matrices || !stage) return c->fail(BBB_EINVAL, "null argument");
    BBB_CUDA(c, cudaSetDevice(c->device));
    c->begin_call();
    c->last_kernel = "rnghunt_eval";
    const uint32_t nw64 = (n + 63) / 64, nwp = 2 * nw64 + 1, pw64 = (n + 1 + 63) / 64;
    const size_t mat_bytes = (size_t)n * nw64 * 8;
    const uint32_t cap = (uint32_t)(ncand < 65536 ? (ncand ? ncand : 1) : 65536);
    HuntShared sh;
    HuntSlot b;
    rc = hunt_setup(c, n, cap, nwp, false, 1, sh, &b);
    if (rc) return rc;
    uint64_t *d_m = (uint64_t *)c->scratch(5, cap * mat_bytes + (size_t)cap * pw64 * 8);
    if (!d_m) return BBB_ECUDA;
    uint64_t *d_poly = (uint64_t *)((char *)d_m + cap * mat_bytes);
    for (uint64_t done = 0; done < ncand; done += cap) {
        const uint32_t nb = (uint32_t)(ncand - done < cap ? ncand - done : cap);
        BBB_CUDA(c, cudaMemsetAsync(b.counters, 0, 16, c->stream));
        BBB_CUDA(c, c->h2d(d_m, matrices + done * n * nw64, nb * mat_bytes));
        HuntOut o;
        o.stage = b.stage; o.surv_count = b.counters; o.surv_cand = b.surv_cand; o.surv_poly = b.surv_poly;
        o.poly_out = poly_words ? d_poly : nullptr;
        const uint32_t blocks = (nb + 63) / 64;
#define DENSE_CASE(N) case N: hunt_front_dense_kernel<N><<<blocks, 64, 0, c->stream>>>(n, nb, d_m, o); break
        switch (nw64) {
            DENSE_CASE(1); DENSE_CASE(2); DENSE_CASE(3); DENSE_CASE(4); DENSE_CASE(5); DENSE_CASE(6);
            DENSE_CASE(7); DENSE_CASE(8);
        default: return c->fail(BBB_EINVAL, "bad n");
        }
#undef DENSE_CASE
        c->last_launches++;
        BBB_CUDA(c, cudaGetLastError());
        rc = hunt_back_end(c, n, nb, nwp, sh, b);
        if (rc) return rc;
        BBB_CUDA(c, c->d2h(stage + done, b.stage, nb));
        if (poly_words) BBB_CUDA(c, c->d2h(poly_words + done * pw64, d_poly, (size_t)nb * pw64 * 8));
        BBB_CUDA(c, cudaStreamSynchronize(c->stream));
        b.nb = 0;
    }
    c->end_call();
    return BBB_OK;
}

}  // extern "C"
