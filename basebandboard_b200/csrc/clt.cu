// clt.cu -- LUTOPT -> CLTGRNG on the device (gateware/bbb/rng.py:21-40, :70-108).
//
// Layout of the computation (DESIGN.md "CLT path"):
//   * the sample range is cut into S independent streams of L = 2^k samples; stream s
//     starts from A^(s*L) . x_level, obtained on the device from host-precomputed
//     jump-ahead operators A^(L*2^b) (seed kernel, one warp per 32 streams);
//   * generators are BIT-SLICED: a thread owns 32 streams, register/shared word i
//     holds state bit i of all 32 (bit j of the word = stream j of the thread);
//   * one LUTOPT clock is the sparse XOR network x'[r] = XOR x[taps(r)] on those
//     words; the CLT tree (rng.py:96-105) equals popcount(x' XOR M) - n/2 with
//     M = indices of odd bit-parity, computed by a carry-save
//     adder network on the same words;
//   * counts are un-sliced with a 32x32 bit transpose and histogrammed / stored.
//
// This file holds the seed kernel, the data-driven kernel that works for ANY
// loaded recurrence ("generic"), and the dispatch; the straight-line kernels
// generated for the shipped matrices live in clt_spec.cu.
#include "bbb_internal.h"
#include "clt_common.cuh"

#include <mutex>

// ---------------------------------------------------------------------------
// seed kernel: planes[i*T + t] = bit i of the 32 streams of main-kernel thread t
// ---------------------------------------------------------------------------
template <int NW32>
__global__ void __launch_bounds__(256)
seed_kernel(const uint32_t *__restrict__ J,   // [nbits][n][NW32] column-major jump operators
            SeedStart xs, uint32_t n, uint32_t nbits, uint32_t T, uint32_t *__restrict__ planes)
{
    const uint32_t gw = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;   // = main-kernel thread
    const uint32_t lane = threadIdx.x & 31;
    if (gw >= T) return;
    const uint32_t s = gw * 32 + lane;   // stream index
    // The warp works on ONE state at a time, replicated in every lane: a jump operator is applied
    // cooperatively (lane l folds columns l, l+32, ... -- coalesced reads of the column-major operator --
    // then a warp-wide XOR reduction (REDUX) leaves the product in all lanes).  First the bits of the stream index the
    // whole warp shares (b >= 5), then 31 successive applications of J[0] = A^L; lane l keeps the state
    // after l of them: A^(s*L) x with s = 32*gw + l.
    uint32_t y[NW32], x[NW32];
#pragma unroll
    for (int w = 0; w < NW32; w++) y[w] = xs.w[w];
    auto apply = [&](const uint32_t *Jb) {
        uint32_t acc[NW32];
#pragma unroll
        for (int w = 0; w < NW32; w++) acc[w] = 0;
#pragma unroll
        for (int k = 0; k < NW32; k++) {
            const uint32_t col = (uint32_t)k * 32u + lane;
            if (col < n && ((y[k] >> lane) & 1u)) {
                const uint32_t *cp = Jb + (size_t)col * NW32;
                if constexpr (NW32 % 4 == 0) {
#pragma unroll
                    for (int w = 0; w < NW32; w += 4) {
                        const uint4 v = __ldg((const uint4 *)(cp + w));
                        acc[w] ^= v.x; acc[w + 1] ^= v.y; acc[w + 2] ^= v.z; acc[w + 3] ^= v.w;
                    }
                } else {
#pragma unroll
                    for (int w = 0; w < NW32; w++) acc[w] ^= __ldg(cp + w);
                }
            }
        }
#pragma unroll
        for (int w = 0; w < NW32; w++) y[w] = __reduce_xor_sync(0xffffffffu, acc[w]);   // REDUX.XOR: one instruction per word
    };
    for (uint32_t b = 5; b < nbits; b++)
        if ((gw >> (b - 5)) & 1u) apply(J + (size_t)b * n * NW32);
#pragma unroll
    for (int w = 0; w < NW32; w++) x[w] = y[w];
    if (nbits > 0) {
        for (uint32_t l = 1; l < 32; l++) {
            apply(J);
            if (lane == l) {
#pragma unroll
                for (int w = 0; w < NW32; w++) x[w] = y[w];
            }
        }
    }
    // ballot transpose: word i of thread gw = bit i of every lane's state
    for (uint32_t i0 = 0; i0 < n; i0 += 32) {
        uint32_t keep = 0;
        uint32_t xw = 0;
#pragma unroll
        for (int w = 0; w < NW32; w++) if ((uint32_t)w == (i0 >> 5)) xw = x[w];
#pragma unroll
        for (int k = 0; k < 32; k++) {
            const uint32_t bal = __ballot_sync(0xffffffffu, (xw >> k) & 1u);
            if (lane == (uint32_t)k) keep = bal;
        }
        if (i0 + lane < n) planes[(size_t)(i0 + lane) * T + gw] = keep;
    }
}

// Four stream groups per warp.  ncu on seed_kernel<8> at 2^19 streams (0.28 ms, 9 % of an 8 GiB word-stream call): the
// L1 data pipe is 90 % busy -- every application of an operator reads all of it (8 KB for n = 256) for ONE state.
// Here a warp carries the states of four consecutive stream groups and a column of the operator, once loaded, is
// folded into all four (mask instead of predicate: acc ^= column & m_v), so the operator traffic per state drops to a
// quarter.  Same arithmetic, same results; used for n = 97..128 and 225..256 (four or eight state words) when there are enough groups to fill the GPU.
template <int NW32>
__global__ void __launch_bounds__(128, 4)
seed4_kernel(const uint32_t *__restrict__ J, SeedStart xs, uint32_t n, uint32_t nbits, uint32_t T,
             uint32_t *__restrict__ planes)
{
    constexpr int V = 4;
    static_assert(NW32 % 4 == 0 && NW32 <= 8, "128-bit operator loads; three sets of V * NW32 words stay in registers");
    const uint32_t gw0 = ((blockIdx.x * blockDim.x + threadIdx.x) >> 5) * V;   // first main-kernel thread of this warp
    const uint32_t lane = threadIdx.x & 31;
    if (gw0 >= T) return;
    // lane l keeps the states after l applications of A^L: parked in shared memory (column = lane), not in 32 registers
    __shared__ uint32_t park[4][V][NW32][32];
    uint32_t (*const xp)[NW32][32] = park[threadIdx.x >> 5];
    uint32_t y[V][NW32];
#pragma unroll
    for (int v = 0; v < V; v++)
#pragma unroll
        for (int w = 0; w < NW32; w++) y[v][w] = xs.w[w];
    auto apply = [&](const uint32_t *Jb, uint32_t vmask) {
        uint32_t acc[V][NW32];
#pragma unroll
        for (int v = 0; v < V; v++)
#pragma unroll
            for (int w = 0; w < NW32; w++) acc[v][w] = 0;
#pragma unroll
        for (int k = 0; k < NW32; k++) {
            const uint32_t col = (uint32_t)k * 32u + lane;
            if (col < n) {
                uint32_t m[V];
#pragma unroll
                for (int v = 0; v < V; v++) m[v] = 0u - ((y[v][k] >> lane) & (vmask >> v) & 1u);
                const uint32_t *cp = Jb + (size_t)col * NW32;
#pragma unroll
                for (int w = 0; w < NW32; w += 4) {
                    const uint4 c4 = __ldg((const uint4 *)(cp + w));
#pragma unroll
                    for (int v = 0; v < V; v++) {
                        acc[v][w] ^= c4.x & m[v]; acc[v][w + 1] ^= c4.y & m[v];
                        acc[v][w + 2] ^= c4.z & m[v]; acc[v][w + 3] ^= c4.w & m[v];
                    }
                }
            }
        }
#pragma unroll
        for (int v = 0; v < V; v++)
            if ((vmask >> v) & 1u) {                      // warp-uniform
#pragma unroll
                for (int w = 0; w < NW32; w++) y[v][w] = __reduce_xor_sync(0xffffffffu, acc[v][w]);
            }
    };
    for (uint32_t b = 5; b < nbits; b++) {
        uint32_t vmask = 0;
#pragma unroll
        for (int v = 0; v < V; v++) vmask |= (((gw0 + v) >> (b - 5)) & 1u) << v;
        if (vmask) apply(J + (size_t)b * n * NW32, vmask);
    }
#pragma unroll
    for (int v = 0; v < V; v++)
#pragma unroll
        for (int w = 0; w < NW32; w++) xp[v][w][lane] = y[v][w];      // nbits == 0: every lane the start state
    if (nbits > 0) {
#pragma unroll 1
        for (uint32_t l = 1; l < 32; l++) {
            apply(J, (1u << V) - 1u);
            if (lane == l) {
#pragma unroll
                for (int v = 0; v < V; v++)
#pragma unroll
                    for (int w = 0; w < NW32; w++) xp[v][w][lane] = y[v][w];
            }
        }
    }
    __syncwarp();
    // ballot transpose: word i of thread gw = bit i of every lane's state
#pragma unroll
    for (int v = 0; v < V; v++) {
        const uint32_t gw = gw0 + v;
        if (gw >= T) break;                               // warp-uniform
#pragma unroll
        for (int w = 0; w < NW32; w++) {
            uint32_t keep = 0;
            const uint32_t xw = xp[v][w][lane];
#pragma unroll
            for (int k = 0; k < 32; k++) {
                const uint32_t bal = __ballot_sync(0xffffffffu, (xw >> k) & 1u);
                if (lane == (uint32_t)k) keep = bal;
            }
            const uint32_t i = (uint32_t)w * 32u + lane;
            if (i < n) planes[(size_t)i * T + gw] = keep;
        }
    }
}

// ---------------------------------------------------------------------------
// generic main kernel: any recurrence, taps read from shared memory
// ---------------------------------------------------------------------------
struct GenericParams {
    uint32_t n, nnz, T, L, model, noise_var, nw32out;
    uint64_t nvalid;            // streams s < nvalid are real
    uint64_t out_base;          // sample offset of stream 0 in the output arrays
    const uint32_t *row_ptr;
    const uint16_t *taps;
    const uint32_t *planes;     // [n][T]
    unsigned long long *hist;   // n+1 bins or null
    int8_t *out8;
    int16_t *out16;
    uint32_t *out_words;        // raw URNG words, nw32out 32-bit words per sample
    uint32_t want_clt;
};

__global__ void __launch_bounds__(256)
clt_generic_kernel(GenericParams p)
{
    extern __shared__ uint32_t smem[];
    const uint32_t B = blockDim.x, tid = threadIdx.x, n = p.n;
    uint32_t *cur = smem;
    uint32_t *nxt = smem + (size_t)n * B;
    uint32_t *sh_hist = nxt + (size_t)n * B;                 // n+1 bins
    uint16_t *sh_rp = (uint16_t *)(sh_hist + n + 1);         // n+1
    uint16_t *sh_tp = sh_rp + (n + 1) + ((n + 1) & 1);       // nnz

    for (uint32_t i = tid; i <= n; i += B) { sh_rp[i] = (uint16_t)p.row_ptr[i]; sh_hist[i] = 0; }
    for (uint32_t i = tid; i < p.nnz; i += B) sh_tp[i] = p.taps[i];
    const uint32_t t = blockIdx.x * B + tid;                 // main thread index
    const bool live = t < p.T;
    for (uint32_t i = 0; i < n; i++) cur[i * B + tid] = live ? p.planes[(size_t)i * p.T + t] : 0u;
    __syncthreads();

    // which of my 32 streams are real
    uint32_t vmask = 0;
    if (live) {
        const uint64_t s0 = (uint64_t)t * 32;
        if (s0 + 32 <= p.nvalid) vmask = 0xffffffffu;
        else if (s0 < p.nvalid) vmask = (1u << (uint32_t)(p.nvalid - s0)) - 1u;
    }
    const uint32_t half = n >> 1;
    // flush the block histogram before a 32-bit bin can overflow
    const uint32_t flush_every = 0x7fffffffu / (B * 32u);
    uint32_t since_flush = 0;

    for (uint32_t step = 0; step < p.L; step++) {
        // ---- one LUTOPT clock: nxt = A . cur (rng.py:38-40) ----
        for (uint32_t r = 0; r < n; r++) {
            uint32_t acc = 0;
            const uint32_t e = sh_rp[r + 1];
            for (uint32_t j = sh_rp[r]; j < e; j++) acc ^= cur[(uint32_t)sh_tp[j] * B + tid];
            nxt[r * B + tid] = acc;
        }
        { uint32_t *tmp = cur; cur = nxt; nxt = tmp; }

        // ---- raw URNG words (LUTOPT.x each clock, rng.py:27) ----
        if (p.out_words) {
            for (uint32_t w = 0; w < p.nw32out; w++) {
                uint32_t a[32];
#pragma unroll
                for (int k = 0; k < 32; k++) {
                    const uint32_t i = w * 32 + k;
                    a[k] = i < n ? cur[i * B + tid] : 0u;
                }
                transpose32(a);
#pragma unroll
                for (int j = 0; j < 32; j++) {
                    if ((vmask >> j) & 1u) {
                        const uint64_t smp = p.out_base + ((uint64_t)t * 32 + j) * p.L + step;
                        p.out_words[smp * p.nw32out + w] = a[j];
                    }
                }
            }
        }

        if (p.want_clt) {
            // ---- popcount(x ^ M): carry-save network, 8 planes per round ----
            uint32_t a1 = 0, a2 = 0, a4 = 0, hi[7] = {0, 0, 0, 0, 0, 0, 0};
            for (uint32_t i0 = 0; i0 < n; i0 += 8) {
                uint32_t q[8];
#pragma unroll
                for (int k = 0; k < 8; k++) {
                    const uint32_t i = i0 + k;
                    // M: inputs of odd index parity enter the tree negated (rng.py:96-105)
                    q[k] = cur[i * B + tid] ^ (0u - (uint32_t)(__popc(i) & 1));
                }
                uint32_t t2a, t2b, t4a, t4b, e8;
                csa(t2a, a1, a1, q[0], q[1]);
                csa(t2b, a1, a1, q[2], q[3]);
                csa(t4a, a2, a2, t2a, t2b);
                csa(t2a, a1, a1, q[4], q[5]);
                csa(t2b, a1, a1, q[6], q[7]);
                csa(t4b, a2, a2, t2a, t2b);
                csa(e8, a4, a4, t4a, t4b);
#pragma unroll
                for (int k = 0; k < 7; k++) {   // ripple the weight-8 carry into bits 3..9
                    const uint32_t c = hi[k] & e8;
                    hi[k] ^= e8;
                    e8 = c;
                }
            }
            uint32_t a[32];
#pragma unroll
            for (int k = 0; k < 32; k++) a[k] = 0;
            a[0] = a1; a[1] = a2; a[2] = a4;
#pragma unroll
            for (int k = 0; k < 7; k++) a[3 + k] = hi[k];
            transpose32(a);   // a[j] = popcount of stream j
#pragma unroll
            for (int j = 0; j < 32; j++) {
                if ((vmask >> j) & 1u) {
                    uint32_t v = a[j];
                    if (p.model == BBB_MODEL_GATEWARE) v &= (n - 1);   // Signal((logn, True)), rng.py:78
                    if (p.hist) atomicAdd(&sh_hist[v], 1u);
                    const int32_t g = (int32_t)v - (int32_t)half;
                    const uint64_t smp = p.out_base + ((uint64_t)t * 32 + j) * p.L + step;
                    if (p.out8) p.out8[smp] = (int8_t)g;
                    if (p.out16) p.out16[smp] = (int16_t)(g * (int32_t)p.noise_var);   // tx.py:75-76
                }
            }
            if (p.hist && ++since_flush >= flush_every) {
                __syncthreads();
                for (uint32_t i = tid; i <= n; i += B) {
                    if (sh_hist[i]) atomicAdd(&p.hist[i], (unsigned long long)sh_hist[i]);
                    sh_hist[i] = 0;
                }
                __syncthreads();
                since_flush = 0;
            }
        }
    }
    if (p.hist) {
        __syncthreads();
        for (uint32_t i = tid; i <= n; i += B)
            if (sh_hist[i]) atomicAdd(&p.hist[i], (unsigned long long)sh_hist[i]);
    }
}

// ---------------------------------------------------------------------------
// host side: geometry, seeding, dispatch
// ---------------------------------------------------------------------------
static int launch_seed(bbb_ctx *c, const uint32_t *dJ, const SeedStart &xs, uint32_t nbits, uint32_t T,
                       uint32_t *d_planes)
{
    const uint32_t n = c->n, nw32 = (n + 31) / 32;
    const uint32_t threads = 256;
    if ((nw32 == 4 || nw32 == 8) && T >= 8u * (uint32_t)c->sm_count && !c->tune_seed_single) {
        const uint32_t warps = (T + 3) / 4, blocks4 = (warps + 3) / 4;       // 128-thread blocks
        if (nw32 == 4) seed4_kernel<4><<<blocks4, 128, 0, c->stream>>>(dJ, xs, n, nbits, T, d_planes);
        else seed4_kernel<8><<<blocks4, 128, 0, c->stream>>>(dJ, xs, n, nbits, T, d_planes);
        c->last_launches++;
        BBB_CUDA(c, cudaGetLastError());
        return BBB_OK;
    }
    const uint32_t blocks = (uint32_t)(((uint64_t)T * 32 + threads - 1) / threads);
#define SEED_CASE(NW) \
    case NW: seed_kernel<NW><<<blocks, threads, 0, c->stream>>>(dJ, xs, n, nbits, T, d_planes); break
    switch (nw32) {
        SEED_CASE(1); SEED_CASE(2); SEED_CASE(3); SEED_CASE(4); SEED_CASE(5); SEED_CASE(6); SEED_CASE(7);
        SEED_CASE(8); SEED_CASE(9); SEED_CASE(10); SEED_CASE(11); SEED_CASE(12); SEED_CASE(13);
        SEED_CASE(14); SEED_CASE(15); SEED_CASE(16);
    default: return c->fail(BBB_EINVAL, "n too large");
    }
#undef SEED_CASE
    c->last_launches++;
    BBB_CUDA(c, cudaGetLastError());
    return BBB_OK;
}

// clt_spec.cu
int bbb_clt_spec_streams_per_block(const bbb_ctx *c, bool hist_only);
int bbb_clt_spec_blocks_per_sm(const bbb_ctx *c, int kind);
uint32_t bbb_clt_spec_wave_blocks(const bbb_ctx *c, int kind);
int bbb_clt_spec_words_streams_per_block(const bbb_ctx *c);
int bbb_clt_spec_launch(bbb_ctx *c, uint32_t blocks, uint32_t L, uint32_t model, const uint32_t *planes,
                        unsigned long long *hist);
int bbb_clt_spec_stream_ok(const bbb_ctx *c, int kind);
int bbb_clt_spec_dec_ok(const bbb_ctx *c);
int bbb_clt_spec_dec_launch(bbb_ctx *c, uint32_t blocks, uint32_t Ld, uint32_t spb, uint32_t delay, const uint32_t *planes,
                            int8_t *out);
int bbb_clt_spec_stream_launch(bbb_ctx *c, int kind, uint32_t blocks, uint32_t L, const uint32_t *planes, void *out,
                               uint32_t noise_var);

// Uploads the jump-ahead operators A^(L*2^b), b < nbits, column-major 32-bit words, into scratch
// slot 0 (kept across calls while (L, nbits) repeat).  Returns the device pointer or null.
static uint32_t *upload_jump_ops(bbb_ctx *c, uint64_t L, uint32_t nbits)
{
    const uint32_t n = c->n, nw32 = (n + 31) / 32;
    const size_t per = (size_t)n * nw32;
    // a few (L, nbits) geometries recur from call to call (a straight-line launch and its tail):
    // keep the most recently used operator sets on the device
    bbb_ctx::JumpSet *slot = &c->jump_sets[0];
    for (auto &js : c->jump_sets) {
        if (js.L == L && js.nbits >= nbits && nbits && js.dev) { js.stamp = ++c->jump_stamp; return js.dev; }
        if (js.stamp < slot->stamp) slot = &js;
    }
    const size_t bytes = (nbits ? nbits : 1) * per * 4;
    if (slot->bytes < bytes) {
        if (slot->dev) cudaFree(slot->dev);
        slot->dev = nullptr;
        slot->bytes = 0;
        if (cudaMalloc(&slot->dev, bytes) != cudaSuccess) { c->cuda_fail(cudaGetLastError(), "cudaMalloc(jump operators)"); return nullptr; }
        slot->bytes = bytes;
    }
    slot->L = 0;
    slot->nbits = 0;
    slot->stamp = ++c->jump_stamp;
    uint32_t *dJ = slot->dev;
    if (!nbits) return dJ;
    std::vector<uint32_t> all(nbits * per), one;
    HostMat M;
    if ((L & (L - 1)) == 0) {
        uint32_t logL = 0;
        while ((1ull << logL) < L) logL++;
        for (uint32_t b = 0; b < nbits; b++) {
            c->pow2(logL + b).to_colmajor32(one);
            memcpy(&all[b * per], one.data(), per * 4);
        }
    } else {
        c->power(L, M);
        for (uint32_t b = 0; b < nbits; b++) {
            M.to_colmajor32(one);
            memcpy(&all[b * per], one.data(), per * 4);
            if (b + 1 < nbits) HostMat::mul(M, M, M);
        }
    }
    if (c->h2d(dJ, all.data(), all.size() * 4) != cudaSuccess ||
        cudaStreamSynchronize(c->stream) != cudaSuccess) {   // `all` is pageable host memory
        c->cuda_fail(cudaGetLastError(), "upload jump operators");
        return nullptr;
    }
    slot->L = L;
    slot->nbits = nbits;
    return dJ;
}

static void advance_state(bbb_ctx *c, uint64_t *x, uint64_t clocks)
{
    for (uint32_t i = 0; i < 64 && (clocks >> i); i++)
        if ((clocks >> i) & 1) c->pow2(i).apply(x, x);
}

static void to_seed(const uint64_t *x, uint32_t nw64, SeedStart &xs)
{
    for (uint32_t w = 0; w < BBB_MAX_NW32; w++)
        xs.w[w] = w < 2 * nw64 ? (uint32_t)(x[w >> 1] >> (32 * (w & 1))) : 0u;
}

int bbb_clt_run_device(bbb_ctx *c, const uint64_t *x0, uint64_t first, uint64_t count, uint32_t model,
                       uint64_t *d_hist, int8_t *d_out8, int16_t *d_out16, uint32_t noise_var,
                       uint32_t *d_out_words32)
{
    const uint32_t n = c->n, nw64 = (n + 63) / 64;
    const bool want_clt = d_hist || d_out8 || d_out16;
    if (want_clt && (n & (n - 1))) return c->fail(BBB_EINVAL, "CLTGRNG needs n a power of two (rng.py:73-74), n=%u", n);
    if (d_out8 && n > 256) return c->fail(BBB_EINVAL, "int8 samples need n <= 256");
    BBB_CUDA(c, cudaSetDevice(c->device));

    // state at `first`
    uint64_t xl[BBB_MAX_NW64] = {0};
    memcpy(xl, x0, nw64 * sizeof(uint64_t));
    if (n % 64) xl[nw64 - 1] &= (1ull << (n % 64)) - 1;
    advance_state(c, xl, first);

    // generic-kernel geometry
    uint32_t Bg = 256;
    while (Bg > 32 && (size_t)8 * n * Bg > 200 * 1024) Bg -= 32;
    const size_t g_smem = (size_t)8 * n * Bg + (n + 1) * 4 + ((n + 2) & ~1u) * 2 + c->taps.size() * 2 + 16;
    const uint64_t Sres = (uint64_t)c->sm_count * Bg * 32;
    const uint32_t Lmin = 64;

    // which straight-line kernel of the shipped n256 matrix applies: 0 fused histogram, 1 int8 samples,
    // 2 raw words, 3 int16 scaled samples, -1 none
    int kind = -1;
    if (d_hist && !d_out8 && !d_out16 && !d_out_words32) kind = 0;
    else if (d_out8 && !d_hist && !d_out16 && !d_out_words32 && ((uintptr_t)d_out8 & 31) == 0) kind = 1;
    else if (d_out_words32 && !d_hist && !d_out8 && !d_out16 && ((uintptr_t)d_out_words32 & 15) == 0) kind = 2;
    else if (d_out16 && !d_hist && !d_out8 && !d_out_words32 && ((uintptr_t)d_out16 & 15) == 0) kind = 3;
    uint64_t spb = 0;
    if (kind == 0) spb = (uint64_t)bbb_clt_spec_streams_per_block(c, true);
    else if (kind == 2 && bbb_clt_spec_stream_ok(c, kind)) spb = (uint64_t)bbb_clt_spec_words_streams_per_block(c);
    else if (kind > 0 && bbb_clt_spec_stream_ok(c, kind)) spb = (uint64_t)bbb_clt_spec_streams_per_block(c, true);

    // The pieces after a straight-line launch (its ragged tail: a seeding kernel + a short data-driven launch, 0.2 ms
    // beside a 3 ms word-stream kernel) are independent of it: they run on ctx->stream3 with their own plane buffer
    // and join the call's stream at the end.
    struct TailFork {
        bbb_ctx *c;
        cudaStream_t main;
        bool forked = false;
        ~TailFork()
        {
            c->stream = main;
            if (forked && cudaEventRecord(c->ev_tail_join, c->stream3) == cudaSuccess)
                cudaStreamWaitEvent(main, c->ev_tail_join, 0);
        }
    } fork{c, c->stream};
    BBB_CUDA(c, cudaEventRecord(c->ev_tail_fork, fork.main));
    uint64_t tail_below = 0;             // streams of the last straight-line launch: what is left after it is < this

    uint64_t done = 0;
    c->last_kernel = "generic";
    while (done < count) {
        const uint64_t rem = count - done;
        if (tail_below && rem < tail_below && !fork.forked) {
            BBB_CUDA(c, cudaStreamWaitEvent(c->stream3, c->ev_tail_fork, 0));
            c->stream = c->stream3;
            fork.forked = true;
        }
        const int planes_slot = fork.forked ? 15 : 1;

        // ---- specialised kernel: `blocks` CTAs x spb streams x L clocks, one wave ----
        if (spb && !fork.forked && ((kind != 1 && kind != 3) || (done & 31) == 0)) {   // (the tail: data-driven only,
                                                                                       // the slabs of the stream kernels are in use)
            const uint64_t max_blocks = c->tune_spec_max_blocks ? c->tune_spec_max_blocks
                                        : (uint64_t)bbb_clt_spec_wave_blocks(c, kind);
            // fewest clocks per stream worth a launch (the stream kernels amortise their seeding sooner)
            const uint64_t min_steps = kind == 2   ? (c->tune_spec_min_steps / 4 > 8 ? c->tune_spec_min_steps / 4 : 8)
                                       : kind > 0 ? (c->tune_spec_min_steps / 4 > 32 ? c->tune_spec_min_steps / 4 : 32)
                                                  : c->tune_spec_min_steps;
            uint64_t blocks = rem / (spb * min_steps);
            if (blocks > max_blocks) blocks = max_blocks;
            uint64_t L = blocks ? rem / (blocks * spb) : 0;
            if (L > (1ull << 26)) L = 1ull << 26;          // private u32 bins: 32 * L < 2^32
            if (kind == 1 || kind == 3) L &= ~31ull;        // 32 clocks per flush (one full sector per stream)
            if (blocks >= 1 && L >= 8) {
                const uint64_t S = blocks * spb;
                const uint32_t T = (uint32_t)(S / 32);
                uint32_t nbits = 0;
                while (nbits < 40 && ((S - 1) >> nbits)) nbits++;
                uint32_t *dJ = upload_jump_ops(c, L, nbits);
                if (!dJ) return BBB_ECUDA;
                uint32_t *d_planes = (uint32_t *)c->scratch(planes_slot, (size_t)n * T * 4);
                if (!d_planes) return BBB_ECUDA;
                // the CLT kernels take the tree of a stream's state BEFORE each clock: start one clock in;
                // the raw-word kernel emits the state AFTER each clock
                uint64_t x1[BBB_MAX_NW64];
                memcpy(x1, xl, sizeof x1);
                if (kind != 2) advance_state(c, x1, 1);
                SeedStart xs;
                to_seed(x1, nw64, xs);
                int rc = launch_seed(c, dJ, xs, nbits, T, d_planes);
                if (rc) return rc;
                if (kind == 0)
                    rc = bbb_clt_spec_launch(c, (uint32_t)blocks, (uint32_t)L, model, d_planes,
                                             (unsigned long long *)d_hist);
                else
                    rc = bbb_clt_spec_stream_launch(c, kind, (uint32_t)blocks, (uint32_t)L, d_planes,
                                                    kind == 1 ? (void *)(d_out8 + done)
                                                    : kind == 3 ? (void *)(d_out16 + done)
                                                                : (void *)(d_out_words32 + done * 8),
                                                    noise_var);
                if (rc) return rc;
                tail_below = S;
                advance_state(c, xl, S * L);
                done += S * L;
                continue;
            }
        }

        // ---- data-driven kernel: S streams of L = 2^k clocks ----
        // short ranges (the ragged tails of the specialised launches): a clock of this kernel costs tens of
        // microseconds whatever the grid, so fill the GPU with streams before adding clocks per stream;
        // long ranges: 64 clocks or more per stream, up to 8 waves of streams
        uint32_t logL = 0;
        while (logL < 6 && (rem >> (logL + 1)) >= Sres / 2) logL++;
        if (logL == 6)
            while (logL < 40 && (rem >> logL) > 8 * Sres) logL++;
        (void)Lmin;
        const uint64_t L = 1ull << logL;
        const uint64_t S = rem >> logL;                     // >= 1
        const uint64_t T64 = (S + 31) / 32;
        if (T64 > 0x7fffffffull / 32) return c->fail(BBB_EINVAL, "range too large");
        const uint32_t T = (uint32_t)T64;
        uint32_t nbits = 0;
        while (nbits < 40 && ((S - 1) >> nbits)) nbits++;
        uint32_t *dJ = upload_jump_ops(c, L, nbits);
        if (!dJ) return BBB_ECUDA;
        uint32_t *d_planes = (uint32_t *)c->scratch(planes_slot, (size_t)n * T * 4);
        if (!d_planes) return BBB_ECUDA;
        SeedStart xs;
        to_seed(xl, nw64, xs);
        int rc = launch_seed(c, dJ, xs, nbits, T, d_planes);
        if (rc) return rc;
        {
            GenericParams p;
            p.n = n; p.nnz = (uint32_t)c->taps.size(); p.T = T; p.L = (uint32_t)L; p.model = model;
            p.noise_var = noise_var; p.nw32out = 2 * nw64; p.nvalid = S; p.out_base = done;
            p.row_ptr = c->d_row_ptr; p.taps = c->d_taps; p.planes = d_planes;
            p.hist = (unsigned long long *)d_hist; p.out8 = d_out8; p.out16 = d_out16;
            p.out_words = d_out_words32; p.want_clt = want_clt ? 1u : 0u;
            if (L > 0xffffffffull) return c->fail(BBB_EINVAL, "L too large");
            BBB_CUDA(c, cudaFuncSetAttribute(clt_generic_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                             (int)g_smem));
            const uint32_t blocks = (T + Bg - 1) / Bg;
            clt_generic_kernel<<<blocks, Bg, g_smem, c->stream>>>(p);
            c->last_launches++;
            BBB_CUDA(c, cudaGetLastError());
        }
        const uint64_t adv = S << logL;
        advance_state(c, xl, adv);
        done += adv;
    }
    return BBB_OK;
}

// out[j] = in[j * spb + delay]
__global__ void decimate_kernel(const int8_t *__restrict__ in, uint64_t nd, uint32_t spb, uint32_t delay, int8_t *__restrict__ out)
{
    for (uint64_t j = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; j < nd; j += (uint64_t)gridDim.x * blockDim.x)
        out[j] = in[j * spb + delay];
}

// Decimated int8 samples, d_out[j] = g[first + j*spb + delay] for j < nd (gateware widths): the sample the
// receiver keeps of each bit period (rx.py:18-19).  The shipped n256 matrix runs clt256_dec_kernel
// (one CLT tree per spb clocks); anything else, and the ragged tail, generates the full stream into a
// bounded scratch buffer and picks every spb-th byte.
int bbb_clt_run_decimated(bbb_ctx *c, const uint64_t *x0, uint64_t first, uint64_t nd, uint32_t spb, uint32_t delay,
                          int8_t *d_out)
{
    const uint32_t n = c->n, nw64 = (n + 63) / 64;
    if (n & (n - 1)) return c->fail(BBB_EINVAL, "CLTGRNG needs n a power of two (rng.py:73-74), n=%u", n);
    if (n > 256) return c->fail(BBB_EINVAL, "int8 samples need n <= 256");
    if (!spb || delay >= spb) return c->fail(BBB_EINVAL, "delay < spb required");
    if (spb == 1) return bbb_clt_run_device(c, x0, first, nd, BBB_MODEL_GATEWARE, nullptr, d_out, nullptr, 1, nullptr);
    BBB_CUDA(c, cudaSetDevice(c->device));
    const bool spec = bbb_clt_spec_dec_ok(c) && !c->tune_dec_full_stream && ((uintptr_t)d_out & 31) == 0;
    const uint64_t spbk = (uint64_t)bbb_clt_spec_streams_per_block(c, true);   // streams per CTA
    uint64_t xl[BBB_MAX_NW64] = {0};
    uint64_t done = 0;
    if (spec) {
        memcpy(xl, x0, nw64 * sizeof(uint64_t));
        advance_state(c, xl, first);
    }
    while (done < nd) {
        const uint64_t rem = nd - done;
        if (spec && spbk && (done & 31) == 0) {
            const uint64_t max_blocks = c->tune_spec_max_blocks ? c->tune_spec_max_blocks
                                        : (uint64_t)c->sm_count * bbb_clt_spec_blocks_per_sm(c, 1);
            uint64_t blocks = rem / (spbk * 32);
            if (blocks > max_blocks) blocks = max_blocks;
            uint64_t Ld = blocks ? rem / (blocks * spbk) : 0;
            if (Ld > (1ull << 31) / spb) Ld = (1ull << 31) / spb;   // clocks per stream (Ld * spb) stay below 2^32
            Ld &= ~31ull;                                   // 32 sampled clocks per flush (one full sector per stream)
            if (blocks >= 1 && Ld >= 32) {
                const uint64_t S = blocks * spbk, L = Ld * spb;
                const uint32_t T = (uint32_t)(S / 32);
                uint32_t nbits = 0;
                while (nbits < 40 && ((S - 1) >> nbits)) nbits++;
                uint32_t *dJ = upload_jump_ops(c, L, nbits);
                if (!dJ) return BBB_ECUDA;
                uint32_t *d_planes = (uint32_t *)c->scratch(1, (size_t)n * T * 4);
                if (!d_planes) return BBB_ECUDA;
                uint64_t x1[BBB_MAX_NW64];                  // the tree of a stream's state BEFORE each clock: start one clock in
                memcpy(x1, xl, sizeof x1);
                advance_state(c, x1, 1);
                SeedStart xs;
                to_seed(x1, nw64, xs);
                int rc = launch_seed(c, dJ, xs, nbits, T, d_planes);
                if (rc) return rc;
                rc = bbb_clt_spec_dec_launch(c, (uint32_t)blocks, (uint32_t)Ld, spb, delay, d_planes, d_out + done);
                if (rc) return rc;
                advance_state(c, xl, S * L);
                done += S * Ld;
                continue;
            }
        }
        // full stream of a bounded piece, then every spb-th byte
        uint64_t piece = rem;
        const uint64_t cap = (1ull << 28) / spb;
        if (piece > cap) piece = cap;
        int8_t *tmp = (int8_t *)c->scratch(13, (size_t)piece * spb);
        if (!tmp) return BBB_ECUDA;
        int rc = bbb_clt_run_device(c, x0, first + done * spb, piece * spb, BBB_MODEL_GATEWARE, nullptr, tmp, nullptr, 1, nullptr);
        if (rc) return rc;
        const uint64_t want = (piece + 255) / 256, cap_blocks = (uint64_t)c->sm_count * 16;
        decimate_kernel<<<(uint32_t)(want < cap_blocks ? want : cap_blocks), 256, 0, c->stream>>>(tmp, piece, spb, delay, d_out + done);
        c->last_launches++;
        BBB_CUDA(c, cudaGetLastError());
        if (spec) advance_state(c, xl, piece * spb);
        done += piece;
    }
    return BBB_OK;
}

// ---------------------------------------------------------------------------
// C ABI
// ---------------------------------------------------------------------------
// ---------------------------------------------------------------------------
// bbb_clt_hist_dev_static: bbb_clt_hist_dev for callers that must not allocate (CUDA graph capture of the work
// around it, tight serving loops, harnesses that forbid cudaMalloc).  No counterpart in the reference.
// All device scratch of the call -- the bit-sliced start states, the jump-ahead
// operators and the zeroed histogram the kernels accumulate into -- lives in ONE file-scope array that every GPU
// loading the library holds until exit (4.51 MiB).  The data-driven kernel only (any loaded recurrence, n <= 512);
// at most BBB_STATIC_MAX_STREAMS streams, so the clocks per stream grow with `count`.
// The array is shared by every call on a device and keeps the last call's contents: the histogram part is cleared
// per call, the rest is overwritten before use; calls that overlap on the GPU (two contexts or streams on one
// device) race.
// ---------------------------------------------------------------------------
#define BBB_STATIC_MAX_T 2048u                                   /* threads of 32 streams */
#define BBB_STATIC_MAX_STREAMS (BBB_STATIC_MAX_T * 32u)
#define BBB_STATIC_MAX_NBITS 16u                                 /* stream index bits */
#define BBB_STATIC_PLANE_WORDS ((size_t)BBB_MAX_N * BBB_STATIC_MAX_T)
#define BBB_STATIC_JUMP_WORDS ((size_t)BBB_STATIC_MAX_NBITS * BBB_MAX_N * BBB_MAX_NW32)
#define BBB_STATIC_HIST_WORDS ((size_t)2048)                     /* (n + 1) uint64 bins, n <= 512 */
#define BBB_STATIC_WORDS (BBB_STATIC_PLANE_WORDS + BBB_STATIC_JUMP_WORDS + BBB_STATIC_HIST_WORDS)
static_assert((BBB_MAX_N + 1) * 2 <= BBB_STATIC_HIST_WORDS, "histogram part too small");
static_assert(BBB_STATIC_WORDS * 4 == BBB_STATIC_SCRATCH_BYTES, "include/bbb_b200.h states the size");
static_assert(BBB_STATIC_PLANE_WORDS % 64 == 0 && BBB_STATIC_JUMP_WORDS % 64 == 0, "parts stay 256-byte aligned");

static __device__ __align__(256) uint32_t bbb_static_scratch[BBB_STATIC_WORDS];

// In host code the array's name is a stand-in, not a device address: the address is asked for once per device, with
// that device current.
static int static_scratch_ptr(bbb_ctx *c, uint32_t **out)
{
    static std::mutex mu;
    static uint32_t *ptrs[64] = {};
    if (c->device < 0 || c->device >= 64) return c->fail(BBB_EINVAL, "device index %d out of range", c->device);
    std::lock_guard<std::mutex> lk(mu);
    if (!ptrs[c->device]) {
        BBB_CUDA(c, cudaSetDevice(c->device));
        void *p = nullptr;
        BBB_CUDA(c, cudaGetSymbolAddress(&p, bbb_static_scratch));
        if (!p) return c->fail(BBB_ECUDA, "cudaGetSymbolAddress returned null");
        ptrs[c->device] = (uint32_t *)p;
    }
    *out = ptrs[c->device];
    return BBB_OK;
}

static int check_loaded(bbb_ctx *c, const uint64_t *x0)
{
    if (!c) return BBB_EINVAL;
    if (!x0) return c->fail(BBB_EINVAL, "x0 is null");
    if (!c->n) return c->fail(BBB_ESTATE, "no recurrence loaded (bbb_lutopt_load)");
    return BBB_OK;
}

extern "C" {

int bbb_clt_hist_dev(bbb_ctx *c, const uint64_t *x0, uint64_t first, uint64_t count, uint32_t model,
                     uint64_t *acc_dev)
{
    int rc = check_loaded(c, x0);
    if (rc) return rc;
    if (!acc_dev) return c->fail(BBB_EINVAL, "acc_dev is null");
    if (model > 1) return c->fail(BBB_EINVAL, "bad model");
    c->begin_call();
    // the kernels accumulate into a zeroed scratch histogram; a tiny kernel then folds it and the
    // moments (exact integer sums over the bins) into acc_dev
    const uint32_t n = c->n;
    uint64_t *d_h = (uint64_t *)c->scratch(2, (n + 1) * sizeof(uint64_t));
    if (!d_h) return BBB_ECUDA;
    BBB_CUDA(c, cudaMemsetAsync(d_h, 0, (n + 1) * sizeof(uint64_t), c->stream));
    rc = bbb_clt_run_device(c, x0, first, count, model, d_h, nullptr, nullptr, 1, nullptr);
    if (rc) return rc;
    fold_hist_kernel<<<1, 512, 0, c->stream>>>((const unsigned long long *)d_h, n,
                                                (unsigned long long *)acc_dev);
    c->last_launches++;
    BBB_CUDA(c, cudaGetLastError());
    c->end_call();
    return BBB_OK;
}

int bbb_clt_hist_dev_static(bbb_ctx *c, const uint64_t *x0, uint64_t first, uint64_t count, uint32_t model,
                            uint64_t *acc_dev)
{
    int rc = check_loaded(c, x0);
    if (rc) return rc;
    if (!acc_dev) return c->fail(BBB_EINVAL, "acc_dev is null");
    if (model > 1) return c->fail(BBB_EINVAL, "bad model");
    const uint32_t n = c->n, nw64 = (n + 63) / 64, nw32 = (n + 31) / 32;
    if (n & (n - 1)) return c->fail(BBB_EINVAL, "CLTGRNG needs n a power of two (rng.py:73-74), n=%u", n);
    if (count > (1ull << BBB_STATIC_MAX_LOG2_COUNT))
        return c->fail(BBB_EINVAL, "bbb_clt_hist_dev_static: count %llu beyond the maximum 2^%u of the no-allocation variant "
                                   "(static scratch of %zu bytes); use bbb_clt_hist_dev",
                       (unsigned long long)count, BBB_STATIC_MAX_LOG2_COUNT, (size_t)BBB_STATIC_WORDS * 4);
    BBB_CUDA(c, cudaSetDevice(c->device));
    uint32_t *base = nullptr;
    rc = static_scratch_ptr(c, &base);
    if (rc) return rc;
    uint32_t *const d_planes = base;
    uint32_t *const d_jump = base + BBB_STATIC_PLANE_WORDS;
    unsigned long long *const d_h = (unsigned long long *)(base + BBB_STATIC_PLANE_WORDS + BBB_STATIC_JUMP_WORDS);
    c->begin_call();
    BBB_CUDA(c, cudaMemsetAsync(d_h, 0, (n + 1) * sizeof(uint64_t), c->stream));   // the kernels ADD into it

    uint64_t xl[BBB_MAX_NW64] = {0};
    memcpy(xl, x0, nw64 * sizeof(uint64_t));
    if (n % 64) xl[nw64 - 1] &= (1ull << (n % 64)) - 1;
    advance_state(c, xl, first);
    uint32_t Bg = 256;
    while (Bg > 32 && (size_t)8 * n * Bg > 200 * 1024) Bg -= 32;
    const size_t g_smem = (size_t)8 * n * Bg + (n + 1) * 4 + ((n + 2) & ~1u) * 2 + c->taps.size() * 2 + 16;
    BBB_CUDA(c, cudaFuncSetAttribute(clt_generic_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)g_smem));
    const size_t per = (size_t)n * nw32;
    std::vector<uint32_t> ops, one;      // host memory only

    uint64_t done = 0;
    while (done < count) {
        // S <= 65536 streams of L = 2^logL clocks; the remainder goes round again with a shorter L
        const uint64_t rem = count - done;
        uint32_t logL = 0;
        while ((rem >> logL) > BBB_STATIC_MAX_STREAMS) logL++;
        const uint64_t S = rem >> logL;
        const uint32_t T = (uint32_t)((S + 31) / 32);
        uint32_t nbits = 0;
        while (nbits < 40 && ((S - 1) >> nbits)) nbits++;
        if (T > BBB_STATIC_MAX_T || nbits > BBB_STATIC_MAX_NBITS) return c->fail(BBB_EINVAL, "static geometry out of range");
        if (nbits) {
            ops.resize((size_t)nbits * per);
            for (uint32_t b = 0; b < nbits; b++) {
                c->pow2(logL + b).to_colmajor32(one);
                memcpy(&ops[b * per], one.data(), per * 4);
            }
            // stream-ordered behind the previous round's kernels; `ops` is pageable: wait before it is rewritten
            BBB_CUDA(c, c->h2d(d_jump, ops.data(), ops.size() * 4));
            BBB_CUDA(c, cudaStreamSynchronize(c->stream));
        }
        SeedStart xs;
        to_seed(xl, nw64, xs);
        rc = launch_seed(c, d_jump, xs, nbits, T, d_planes);
        if (rc) return rc;
        GenericParams p;
        p.n = n; p.nnz = (uint32_t)c->taps.size(); p.T = T; p.L = 1u << logL; p.model = model;
        p.noise_var = 1; p.nw32out = 2 * nw64; p.nvalid = S; p.out_base = done;
        p.row_ptr = c->d_row_ptr; p.taps = c->d_taps; p.planes = d_planes;
        p.hist = d_h; p.out8 = nullptr; p.out16 = nullptr; p.out_words = nullptr; p.want_clt = 1u;
        clt_generic_kernel<<<(T + Bg - 1) / Bg, Bg, g_smem, c->stream>>>(p);
        c->last_launches++;
        BBB_CUDA(c, cudaGetLastError());
        advance_state(c, xl, S << logL);
        done += S << logL;
    }
    fold_hist_kernel<<<1, 512, 0, c->stream>>>(d_h, n, (unsigned long long *)acc_dev);
    c->last_launches++;
    BBB_CUDA(c, cudaGetLastError());
    c->last_kernel = "generic_static";
    c->end_call();
    return BBB_OK;
}

int bbb_clt_hist(bbb_ctx *c, const uint64_t *x0, uint64_t first, uint64_t count, uint32_t model,
                 uint64_t *hist, int64_t *sum, uint64_t *sumsq)
{
    int rc = check_loaded(c, x0);
    if (rc) return rc;
    const uint32_t n = c->n;
    uint64_t *d_acc = (uint64_t *)c->scratch(3, (n + 3) * sizeof(uint64_t));
    if (!d_acc) return BBB_ECUDA;
    BBB_CUDA(c, cudaMemsetAsync(d_acc, 0, (n + 3) * sizeof(uint64_t), c->stream));
    rc = bbb_clt_hist_dev(c, x0, first, count, model, d_acc);
    if (rc) return rc;
    uint64_t h[BBB_MAX_N + 3];
    BBB_CUDA(c, c->d2h(h, d_acc, (n + 3) * sizeof(uint64_t)));
    BBB_CUDA(c, cudaStreamSynchronize(c->stream));
    const uint32_t bins = n + (model == BBB_MODEL_NUMPY ? 1 : 0);
    if (hist) for (uint32_t b = 0; b < bins; b++) hist[b] += h[b];
    if (sum) *sum += (int64_t)h[n + 1];
    if (sumsq) *sumsq += h[n + 2];
    return BBB_OK;
}

int bbb_clt_stream_dev(bbb_ctx *c, const uint64_t *x0, uint64_t first, uint64_t count, uint32_t model,
                       uint32_t noise_var, int8_t *out8_dev, int16_t *out16_dev)
{
    int rc = check_loaded(c, x0);
    if (rc) return rc;
    if ((out8_dev != nullptr) == (out16_dev != nullptr))
        return c->fail(BBB_EINVAL, "exactly one of out8/out16 must be given");
    if (model > 1) return c->fail(BBB_EINVAL, "bad model");
    if (noise_var > 15 && out16_dev && (uint64_t)noise_var * (c->n / 2) > 32767)
        return c->fail(BBB_EINVAL, "noise_var too large for int16");
    c->begin_call();
    rc = bbb_clt_run_device(c, x0, first, count, model, nullptr, out8_dev, out16_dev, noise_var, nullptr);
    if (rc) return rc;
    c->end_call();
    return BBB_OK;
}

int bbb_clt_stream(bbb_ctx *c, const uint64_t *x0, uint64_t first, uint64_t count, uint32_t model,
                   uint32_t noise_var, int8_t *out8, int16_t *out16)
{
    int rc = check_loaded(c, x0);
    if (rc) return rc;
    if ((out8 != nullptr) == (out16 != nullptr))
        return c->fail(BBB_EINVAL, "exactly one of out8/out16 must be given");
    const size_t eb = out8 ? 1 : 2;
    // 2^28-byte chunks: each chunk re-seeds its streams by jump-ahead, which must stay small beside the chunk
    return bbb_stream_to_host(c, 4, out8 ? (void *)out8 : (void *)out16, (size_t)count * eb, 65536, (size_t)256 << 20,
                              [&](size_t off, size_t bytes, void *dev) {
                                  return bbb_clt_stream_dev(c, x0, first + off / eb, bytes / eb, model, noise_var,
                                                            out8 ? (int8_t *)dev : nullptr, out16 ? (int16_t *)dev : nullptr);
                              });
}

int bbb_clt_stream_dec_dev(bbb_ctx *c, const uint64_t *x0, uint64_t first, uint64_t count, uint32_t spb, uint32_t delay,
                           int8_t *out8_dev)
{
    int rc = check_loaded(c, x0);
    if (rc) return rc;
    if (!out8_dev) return c->fail(BBB_EINVAL, "out is null");
    if (!spb || spb > 64 || delay >= spb) return c->fail(BBB_EINVAL, "1 <= spb <= 64 and delay < spb (rx.py:18-19)");
    if ((first + count * spb) >> 62) return c->fail(BBB_EINVAL, "sample range too large");
    c->begin_call();
    rc = bbb_clt_run_decimated(c, x0, first, count, spb, delay, out8_dev);
    if (rc) return rc;
    c->end_call();
    return BBB_OK;
}

int bbb_clt_stream_dec(bbb_ctx *c, const uint64_t *x0, uint64_t first, uint64_t count, uint32_t spb, uint32_t delay,
                       int8_t *out8)
{
    int rc = check_loaded(c, x0);
    if (rc) return rc;
    if (!out8) return c->fail(BBB_EINVAL, "out is null");
    return bbb_stream_to_host(c, 4, out8, (size_t)count, 65536, (size_t)256 << 20,
                              [&](size_t off, size_t bytes, void *dev) {
                                  return bbb_clt_stream_dec_dev(c, x0, first + (uint64_t)off * spb, bytes, spb, delay,
                                                                (int8_t *)dev);
                              });
}

int bbb_lutopt_stream_dev(bbb_ctx *c, const uint64_t *x0, uint64_t first, uint64_t count,
                          uint64_t *out_words_dev)
{
    int rc = check_loaded(c, x0);
    if (rc) return rc;
    if (!out_words_dev) return c->fail(BBB_EINVAL, "out is null");
    c->begin_call();
    rc = bbb_clt_run_device(c, x0, first, count, 0, nullptr, nullptr, nullptr, 1, (uint32_t *)out_words_dev);
    if (rc) return rc;
    c->end_call();
    return BBB_OK;
}

int bbb_lutopt_stream(bbb_ctx *c, const uint64_t *x0, uint64_t first, uint64_t count, uint64_t *out_words)
{
    int rc = check_loaded(c, x0);
    if (rc) return rc;
    if (!out_words) return c->fail(BBB_EINVAL, "out is null");
    const size_t step_bytes = (size_t)((c->n + 63) / 64) * 8;
    return bbb_stream_to_host(c, 4, out_words, (size_t)count * step_bytes, step_bytes * 4096, (size_t)256 << 20,
                              [&](size_t off, size_t bytes, void *dev) {
                                  return bbb_lutopt_stream_dev(c, x0, first + off / step_bytes, bytes / step_bytes,
                                                               (uint64_t *)dev);
                              });
}

}  // extern "C"
