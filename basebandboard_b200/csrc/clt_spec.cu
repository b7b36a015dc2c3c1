// clt_spec.cu -- straight-line kernel specialised to the shipped n256 recurrence
// (gateware/bbb/rng_recurrences.py:172-259, the matrix tx.py:70 instantiates): fused
// LUTOPT clock + CLTGRNG popcount tree + histogram, 32 bit-sliced streams per thread.
//
// Data layout (DESIGN.md "clt256 kernel"):
//   * 128 threads per CTA, one CTA per SM; a thread's 256 state planes live in shared memory as
//     uint4 groups, word (G*128 + tid): conflict-free LDS.128 / STS.128.  Three regions of 32
//     groups: RA and RB alternate as "old" / "new" copy of storage positions 0..127, R2 holds
//     positions 128..255 and is updated in place at the end of pass 2 (1.5 buffers = 1.5 KB/thread);
//   * storage order is the permutation tools/gen_clt256.py chose (CLT256_PERM);
//   * histogram: per-lane private 64-bin u32 columns in shared memory (bin b of lane l of warp w
//     at word w*2048 + b*32 + l), updated by plain conflict-free read-modify-writes that the
//     generator interleaves with the next clock's instructions; samples outside
//     96 <= popcount < 160 (|g| >= 32, about 6e-5 of them) take a rare atomic slow path into a
//     257-bin CTA histogram.  No barrier inside the clock loop: a thread touches only its own words.
#include "bbb_internal.h"
#include "clt_common.cuh"

#define CLT256_B 128
#define CLT256_STATE_BYTES (96 * CLT256_B * 16)
#define CLT256_FAST_BYTES (4 * 64 * 32 * 4)
#define CLT256_SLOW_BINS 257
#define CLT256_SMEM (CLT256_STATE_BYTES + CLT256_FAST_BYTES + 1040)

#define CLT256_SECTION_PERM
#include "clt256_gen.inc"
#undef CLT256_SECTION_PERM

// popcount outside [96,160): exact slow path.  `outm` = streams of this thread that are out of range.
static __device__ __noinline__ void clt256_slow(uint32_t p0, uint32_t p1, uint32_t p2, uint32_t p3, uint32_t p4,
                                                uint32_t p5, uint32_t p6, uint32_t p7, uint32_t p8, uint32_t outm,
                                                uint32_t model, uint32_t *slow, char *hpb)
{
    while (outm) {
        const uint32_t j = __ffs(outm) - 1;
        outm &= outm - 1;
        uint32_t v = ((p0 >> j) & 1u) | (((p1 >> j) & 1u) << 1) | (((p2 >> j) & 1u) << 2) | (((p3 >> j) & 1u) << 3) |
                     (((p4 >> j) & 1u) << 4) | (((p5 >> j) & 1u) << 5) | (((p6 >> j) & 1u) << 6) |
                     (((p7 >> j) & 1u) << 7) | (((p8 >> j) & 1u) << 8);
        if (model == BBB_MODEL_GATEWARE) v &= 255u;   // Signal((logn, True)), rng.py:78
        atomicAdd(&slow[v], 1u);
        // the masked index planes send this stream to private bin 0 later on: cancel that count
        *(uint32_t *)hpb -= 1u;
    }
}

#define CLT256_AFTER_COUNT(p0, p1, p2, p3, p4, p5, p6, p7, p8, inr) \
    if ((inr) != 0xffffffffu) clt256_slow(p0, p1, p2, p3, p4, p5, p6, p7, p8, ~(inr), model, slow, hpb);

__global__ void __launch_bounds__(CLT256_B, 1)
clt256_hist_kernel(const uint32_t *__restrict__ planes, uint32_t T, uint32_t L, uint32_t model,
                   unsigned long long *__restrict__ hist)
{
    extern __shared__ uint4 smem4[];
    const uint32_t tid = threadIdx.x;
    const uint32_t t = blockIdx.x * CLT256_B + tid;
    char *smb = (char *)smem4;
    uint32_t *fast = (uint32_t *)(smb + CLT256_STATE_BYTES);
    uint32_t *slow = (uint32_t *)(smb + CLT256_STATE_BYTES + CLT256_FAST_BYTES);
    for (uint32_t i = tid; i < CLT256_FAST_BYTES / 4; i += CLT256_B) fast[i] = 0;
    for (uint32_t i = tid; i < CLT256_SLOW_BINS; i += CLT256_B) slow[i] = 0;
    {
        uint32_t *sw = (uint32_t *)smem4;
        for (uint32_t p = 0; p < 256; p++) {
            const uint32_t v = planes[(size_t)CLT256_PERM[p] * T + t];
            const uint32_t G = p < 128 ? (p >> 2) : 64 + ((p - 128) >> 2);
            sw[(G * CLT256_B + tid) * 4 + (p & 3)] = v;
        }
    }
    __syncthreads();
    char *hpb = (char *)(fast + (tid >> 5) * 2048 + (tid & 31));
    uint4 *lo_old = smem4 + tid;
    uint4 *lo_new = smem4 + 32 * CLT256_B + tid;
    uint4 *const hi = smem4 + 64 * CLT256_B + tid;
    uint32_t z0 = 0, z1 = 0, z2 = 0, z3 = 0, z4 = 0, z5 = 0, z6 = 0, z7 = 0;

#pragma unroll 1
    for (uint32_t it = 0; it < L; it++) {
#define CLT256_SECTION_BODY
#include "clt256_gen.inc"
#undef CLT256_SECTION_BODY
        uint4 *tmp = lo_old;
        lo_old = lo_new;
        lo_new = tmp;
    }
    // the last clock's second half of read-modify-writes (words z4..z7)
    {
        const uint32_t zz[4] = {z4, z5, z6, z7};
#pragma unroll
        for (int k = 0; k < 4; k++)
#pragma unroll
            for (int m = 0; m < 4; m++) *(uint32_t *)(hpb + (((zz[k] >> (8 * m)) & 0x3fu) << 7)) += 1u;
    }
    *(uint32_t *)hpb -= 16u;   // the first clock's pass 1 ran 16 updates on the zero-initialised z4..z7
    __syncthreads();
    if (tid < 64) {
        unsigned long long s = 0;
        for (uint32_t w = 0; w < 4; w++)
            for (uint32_t l = 0; l < 32; l++) s += fast[w * 2048 + tid * 32 + ((l + tid) & 31)];
        if (s) atomicAdd(&hist[96 + tid], s);
    }
    for (uint32_t i = tid; i < CLT256_SLOW_BINS; i += CLT256_B)
        if (slow[i]) atomicAdd(&hist[i], (unsigned long long)slow[i]);
}

// Streams one CTA carries, and whether the loaded recurrence has a specialised kernel.
int bbb_clt_spec_streams_per_block(const bbb_ctx *c, bool hist_only)
{
    if (!hist_only || c->shipped != 256 || c->tune_no_spec) return 0;
    return CLT256_B * 32;
}

// Launches the fused kernel on `blocks` CTAs; planes = [256][T] bit-sliced states whose CLT value is
// the first sample of each stream, T = blocks * 128.
int bbb_clt_spec_launch(bbb_ctx *c, uint32_t blocks, uint32_t L, uint32_t model, const uint32_t *planes,
                        unsigned long long *hist)
{
    static_assert(CLT256_SMEM <= 227 * 1024, "shared memory budget");
    BBB_CUDA(c, cudaFuncSetAttribute(clt256_hist_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, CLT256_SMEM));
    clt256_hist_kernel<<<blocks, CLT256_B, CLT256_SMEM, c->stream>>>(planes, blocks * CLT256_B, L, model, hist);
    c->last_launches++;
    BBB_CUDA(c, cudaGetLastError());
    c->last_kernel = "clt256_fused";
    return BBB_OK;
}
