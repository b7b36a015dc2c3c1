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
//
// Kernels in this file (PRODUCT = what the default build launches; the others stay reachable through tuning knobs so
// that the GPU tests keep them bit-exact and DESIGN.md's "measured and rejected" numbers can be reproduced):
//   clt256_hist2_kernel   PRODUCT  fused generate + histogram, one pass per clock (clt256_v2_gen.inc)
//   clt256_hist_kernel    knob clt_spec_variant = 1: the round-1 two-pass kernel (clt256_gen.inc, layout above)
//   clt256_s8v2_kernel    PRODUCT  int8 / int16 sample streams;  clt256_s8_kernel: two-pass predecessor
//   clt256_dec_kernel     PRODUCT  one sample per bit period (slicer loop)
//   clt256_words7_kernel  PRODUCT  raw URNG words, stage in tensor memory, producer / consumer warps
//   clt256_words_kernel, _words2 ... _words6, _words8 ... _words10: earlier and alternative raw-word kernels
//                         (knobs clt_spec_variant = 1, words_variant = 2 ... 16, words_b192)
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

// ===========================================================================
// one-pass variant (tools/gen_clt256_v2.py): the old state is read once per clock, rows are folded
// as their taps arrive, new planes are stored in place.  State: slot p of thread tid at word
// p*128 + tid (conflict-free LDS.32/STS.32), 1.0 buffer = 128 KB per CTA.  The private histogram is
// reached through explicit shared-space addresses in volatile asm: ordered among themselves, and
// invisible to the compiler's alias analysis of the state words (which sees constant offsets only).
// ===========================================================================
#define CLT256V2_B 128
#define CLT256V2_STATE_BYTES (CLT256V2_NSLOTS * CLT256V2_B * 4)
#define CLT256V2_SLOW_WORDS 264                       // 257 bins + [257] = slow-path samples (cancel bin 0)
#define CLT256V2_SMEM (CLT256V2_STATE_BYTES + CLT256_FAST_BYTES + CLT256V2_SLOW_WORDS * 4)
#define CLT256V2_CTAS_PER_SM 2                        // 2 x (state + bins + 1 KB) <= 228 KB, 2 x 128 x 254 registers

// (a & m) | (b & ~m), m a literal: one LOP3 (clt_common.cuh)
#define CLT256V2_BSEL(m, a, b) bit_select<m>((a), (b))
#define CLT256V2_SECTION_PERM
#include "clt256_v2_gen.inc"
#undef CLT256V2_SECTION_PERM

static __device__ __noinline__ void clt256v2_slow(uint32_t p0, uint32_t p1, uint32_t p2, uint32_t p3, uint32_t p4,
                                                  uint32_t p5, uint32_t p6, uint32_t p7, uint32_t p8, uint32_t outm,
                                                  uint32_t model, uint32_t *slow)
{
    while (outm) {
        const uint32_t j = __ffs(outm) - 1;
        outm &= outm - 1;
        uint32_t v = ((p0 >> j) & 1u) | (((p1 >> j) & 1u) << 1) | (((p2 >> j) & 1u) << 2) | (((p3 >> j) & 1u) << 3) |
                     (((p4 >> j) & 1u) << 4) | (((p5 >> j) & 1u) << 5) | (((p6 >> j) & 1u) << 6) |
                     (((p7 >> j) & 1u) << 7) | (((p8 >> j) & 1u) << 8);
        if (model == BBB_MODEL_GATEWARE) v &= 255u;   // Signal((logn, True)), rng.py:78
        atomicAdd(&slow[v], 1u);
        atomicAdd(&slow[257], 1u);                    // its masked index lands in private bin 0: cancelled at the end
    }
}

#define CLT256V2_AFTER_COUNT(p0, p1, p2, p3, p4, p5, p6, p7, p8, inr) \
    if ((inr) != 0xffffffffu) clt256v2_slow(p0, p1, p2, p3, p4, p5, p6, p7, p8, ~(inr), model, slow);

// byte m of z = 6-bit bin of one sample; bin b of this thread's column is 128 bytes * b above hp.
// CLT256V2_FMA_EXTRACT: the byte is isolated on the FMA pipe (z * 2^(24-8m), then the high word of
// * 2^8, multipliers read from registers ptxas cannot fold) instead of one PRMT on the ALU pipe.
#ifndef CLT256V2_FMA_EXTRACT
#define CLT256V2_FMA_EXTRACT 0   /* measured: 459 vs 471 Gsamples/s -- the multiply-high chain costs more than the PRMT it saves */
#endif
#if CLT256V2_FMA_EXTRACT
#define CLT256V2_RMW_LD(nm, z, m)                                                                         \
    uint32_t a##nm, h##nm;                                                                                \
    asm volatile("{\n\t.reg .b32 s, t;\n\tmul.lo.u32 s, %2, %4;\n\tmul.hi.u32 t, s, %5;\n\tmad.lo.u32 %0, t, 128, %3;\n\t" \
                 "ld.shared.u32 %1, [%0];\n\t}"                                                           \
                 : "=r"(a##nm), "=r"(h##nm) : "r"(z), "r"(hp), "r"(kshl[m]), "r"(k256));
#else
#define CLT256V2_RMW_LD(nm, z, m)                                                                        \
    uint32_t a##nm, h##nm;                                                                               \
    asm volatile("{\n\t.reg .b32 t;\n\tprmt.b32 t, %2, 0, 0x444" #m ";\n\tmad.lo.u32 %0, t, 128, %3;\n\t" \
                 "ld.shared.u32 %1, [%0];\n\t}"                                                          \
                 : "=r"(a##nm), "=r"(h##nm) : "r"(z), "r"(hp));
#endif
#define CLT256V2_RMW_ST(nm)                                                                              \
    asm volatile("{\n\t.reg .b32 t;\n\tmad.lo.u32 t, %1, %2, %2;\n\tst.shared.u32 [%0], t;\n\t}" ::"r"(a##nm), "r"(h##nm), "r"(one));

__global__ void __launch_bounds__(CLT256V2_B, CLT256V2_CTAS_PER_SM)
clt256_hist2_kernel(const uint32_t *__restrict__ planes, uint32_t T, uint32_t L, uint32_t model, uint32_t one,
                    unsigned long long *__restrict__ hist)
{
    extern __shared__ uint4 smem4[];
    const uint32_t tid = threadIdx.x;
    const uint32_t t = blockIdx.x * CLT256V2_B + tid;
    char *smb = (char *)smem4;
    uint32_t *fast = (uint32_t *)(smb + CLT256V2_STATE_BYTES);
    uint32_t *slow = (uint32_t *)(smb + CLT256V2_STATE_BYTES + CLT256_FAST_BYTES);
    for (uint32_t i = tid; i < CLT256_FAST_BYTES / 4; i += CLT256V2_B) fast[i] = 0;
    for (uint32_t i = tid; i < CLT256V2_SLOW_WORDS; i += CLT256V2_B) slow[i] = 0;
    uint32_t *const st = (uint32_t *)smem4 + tid;
    for (uint32_t p = 0; p < CLT256V2_NSLOTS; p++) st[p * CLT256V2_B] = planes[(size_t)CLT256V2_SLOT_BIT[p] * T + t];
    __syncthreads();
    const uint32_t hp = (uint32_t)__cvta_generic_to_shared(fast + (tid >> 5) * 2048 + (tid & 31));
    const uint32_t k256 = one << 8;                                  // `one` is 1, unknown to the compiler
    const uint32_t kshl[4] = {one << 24, one << 16, one << 8, one};
    (void)k256; (void)kshl;
    const uint32_t kshr1 = one << 31, kshr2 = one << 30, kshr4 = one << 28;   // x >> s == umulhi(x, 2^(32-s))
#ifndef CLT256V2_FMA_SHR
#define CLT256V2_FMA_SHR 1
#endif
#if CLT256V2_FMA_SHR
#define CLT256V2_SHR(x, s) __umulhi((x), kshr##s)
#else
#define CLT256V2_SHR(x, s) ((x) >> (s))
#endif
    (void)kshr1; (void)kshr2; (void)kshr4;
#define CLT256V2_SECTION_RESIDENT_DECL
#include "clt256_v2_gen.inc"
#undef CLT256V2_SECTION_RESIDENT_DECL

    // Software-pipelined: iteration `it` reads the state before clock `it`, feeds it to the popcount
    // network and clocks the recurrence; the last levels of that network, the range test and the
    // un-slicing of clock it-1 (carried in the k<w>_<i> planes) are interleaved with its first steps.
#pragma unroll 1
    for (uint32_t it = 0; it < L; it++) {
#define CLT256V2_SECTION_BODY
#include "clt256_v2_gen.inc"
#undef CLT256V2_SECTION_BODY
    }
    __syncthreads();
    {
        // the last clock's samples; the carried planes started as 32 fake samples of popcount 128 (bin 32)
#define CLT256V2_SECTION_TAIL
#include "clt256_v2_gen.inc"
#undef CLT256V2_SECTION_TAIL
        uint32_t *col = fast + (tid >> 5) * 2048 + (tid & 31);
        const uint32_t zz[8] = {z0, z1, z2, z3, z4, z5, z6, z7};
#pragma unroll
        for (int k = 0; k < 8; k++)
#pragma unroll
            for (int m = 0; m < 4; m++) col[((zz[k] >> (8 * m)) & 0x3fu) * 32] += 1u;
        col[32 * 32] -= 32u;
    }
    __syncthreads();
    if (tid < 64) {
        unsigned long long s = 0;
        for (uint32_t w = 0; w < 4; w++)
            for (uint32_t l = 0; l < 32; l++) s += fast[w * 2048 + tid * 32 + ((l + tid) & 31)];
        if (tid == 0) s -= slow[257];
        if (s) atomicAdd(&hist[96 + tid], s);
    }
    for (uint32_t i = tid; i < 257; i += CLT256V2_B)
        if (slow[i]) atomicAdd(&hist[i], (unsigned long long)slow[i]);
}

// CTAs of the fused kernel one SM holds at a time
int bbb_clt_spec_blocks_per_sm(const bbb_ctx *c, int kind)
{
    if (kind == 2) return (c->tune_spec_variant != 1 && (c->tune_words_variant == 4 || c->tune_words_variant == 5)) ? 2 : 1;
    return c->tune_spec_variant == 1 ? 1 : CLT256V2_CTAS_PER_SM;
}

// CTAs of one wave of a straight-line launch.  The raw-word kernel (clt256_words7_kernel) is the exception to "fill
// every SM": each CTA keeps 4096 streams open and a stream's 128-byte line stays in L2 for four clocks, so the open
// lines of 148 CTAs (78 MB) no longer fit beside each other and leave as single sectors.  The best count moves down
// as the per-SM rate goes up: 128 CTAs while the compute warps did the whole transpose (2859 against 2255 GB/s at
// 148), 112 = three quarters of the SMs (rounded up to a multiple of 8) since the drain warps do two of its five
// stages (3200 GB/s on two boxes; 128: 2810-3040, 104: 2980, 120: 2990; profiles/r2_probe_words7.txt).
uint32_t bbb_clt_spec_wave_blocks(const bbb_ctx *c, int kind)
{
    const uint32_t sms = (uint32_t)c->sm_count;
    const bool words7 = kind == 2 && c->tune_spec_variant != 1 && !c->tune_words_b192 &&
                        (c->tune_words_variant == 0 || c->tune_words_variant >= 7);
    if (!words7) return sms * (uint32_t)bbb_clt_spec_blocks_per_sm(c, kind);
    if (c->tune_words_variant == 7 || c->tune_words_variant == 8) {      // the whole transpose on the compute warps
        uint32_t p2 = 1;
        while (p2 * 2 <= sms) p2 *= 2;
        return p2 * 4 >= sms * 3 ? p2 : sms * 7 / 8;
    }
    const uint32_t w = ((sms * 3 + 3) / 4 + 7) & ~7u;
    return w < sms ? w : sms;
}

// Streams one CTA carries, and whether the loaded recurrence has a specialised kernel.
int bbb_clt_spec_streams_per_block(const bbb_ctx *c, bool hist_only)
{
    if (!hist_only || c->shipped != 256 || c->tune_no_spec) return 0;
    return CLT256_B * 32;
}

// the raw-word kernel runs 128-thread CTAs; the knob "words_b192" selects the 192-thread experiment (clt256_words3_kernel)
int bbb_clt_spec_words_streams_per_block(const bbb_ctx *c)
{
    if (c->shipped != 256 || c->tune_no_spec) return 0;
    return (c->tune_spec_variant != 1 && c->tune_words_b192) ? 192 * 32 : CLT256_B * 32;
}

// Launches the fused kernel on `blocks` CTAs; planes = [256][T] bit-sliced states whose CLT value is
// the first sample of each stream, T = blocks * 128.
int bbb_clt_spec_launch(bbb_ctx *c, uint32_t blocks, uint32_t L, uint32_t model, const uint32_t *planes,
                        unsigned long long *hist)
{
    static_assert(CLT256_SMEM <= 227 * 1024, "shared memory budget");
    BBB_CUDA(c, cudaFuncSetAttribute(clt256_hist_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, CLT256_SMEM));
    if (c->tune_spec_variant == 1) {
        clt256_hist_kernel<<<blocks, CLT256_B, CLT256_SMEM, c->stream>>>(planes, blocks * CLT256_B, L, model, hist);
        c->last_kernel = "clt256_fused";
    } else {
        static_assert(CLT256V2_CTAS_PER_SM * (CLT256V2_SMEM + 1024) <= 228 * 1024, "shared memory budget");
        BBB_CUDA(c, cudaFuncSetAttribute(clt256_hist2_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, CLT256V2_SMEM));
        // `one`: the +1 of a bin update as h * one + one keeps it on the FMA pipe (the ALU pipe is the bound)
        clt256_hist2_kernel<<<blocks, CLT256V2_B, CLT256V2_SMEM, c->stream>>>(planes, blocks * CLT256V2_B, L, model, 1u, hist);
        c->last_kernel = "clt256_fused2";
    }
    c->last_launches++;
    BBB_CUDA(c, cudaGetLastError());
    return BBB_OK;
}

// ===========================================================================
// materialised streams of the shipped n256 recurrence
// ===========================================================================
#define CLT256S8_STAGE_BYTES (16 * CLT256_B * 16)
#define CLT256S8_SMEM (CLT256_STATE_BYTES + CLT256S8_STAGE_BYTES)

// int8 samples g[i] (gateware width): 8 clocks are staged in shared memory as un-sliced bytes
// (word s of clock tau: byte m = sample of stream 8m+s), then each stream's 8 bytes leave as one
// 64-bit store; sectors are completed in L2 by the following flushes.
#define CLT256_S8_STAGE(z0, z1, z2, z3, z4, z5, z6, z7, pc8)                  \
    stage[((it & 7u) * 2u) * CLT256_B] = make_uint4(z0, z1, z2, z3);          \
    stage[((it & 7u) * 2u + 1u) * CLT256_B] = make_uint4(z4, z5, z6, z7);

__device__ __forceinline__ void bytes4x4(uint32_t a, uint32_t b, uint32_t c, uint32_t d, uint32_t (&o)[4])
{
    const uint32_t ab_lo = __byte_perm(a, b, 0x5140), ab_hi = __byte_perm(a, b, 0x7362);
    const uint32_t cd_lo = __byte_perm(c, d, 0x5140), cd_hi = __byte_perm(c, d, 0x7362);
    o[0] = __byte_perm(ab_lo, cd_lo, 0x5410);
    o[1] = __byte_perm(ab_lo, cd_lo, 0x7632);
    o[2] = __byte_perm(ab_hi, cd_hi, 0x5410);
    o[3] = __byte_perm(ab_hi, cd_hi, 0x7632);
}

// four int8 samples of one word -> four int16 samples times noise_var (tx.py:75-76)
__device__ __forceinline__ int32_t sext_byte(uint32_t x, uint32_t sel)
{
    uint32_t d;   // prmt with the selector's bit 3 set replicates the sign of the chosen byte
    asm("prmt.b32 %0, %1, %2, %3;" : "=r"(d) : "r"(x), "r"(0u), "r"(sel));
    return (int32_t)d;
}

__device__ __forceinline__ uint2 scale4(uint32_t x, int32_t nv)
{
    const int32_t s0 = sext_byte(x, 0x8880u) * nv, s1 = sext_byte(x, 0x9991u) * nv;
    const int32_t s2 = sext_byte(x, 0xAAA2u) * nv, s3 = sext_byte(x, 0xBBB3u) * nv;
    return make_uint2(__byte_perm((uint32_t)s0, (uint32_t)s1, 0x5410), __byte_perm((uint32_t)s2, (uint32_t)s3, 0x5410));
}

// noise_var < 0: out is int8 g[i]; otherwise out is int16 g[i] * noise_var.  (A run-time flag, not a
// template parameter: cudafe++ needs tens of minutes to instantiate a template around the generated body.)
__global__ void __launch_bounds__(CLT256_B, 1)
clt256_s8_kernel(const uint32_t *__restrict__ planes, uint32_t T, uint32_t L, int8_t *__restrict__ out, int32_t noise_var)
{
    const bool OUT16 = noise_var >= 0;
    extern __shared__ uint4 smem4[];
    const uint32_t tid = threadIdx.x;
    const uint32_t t = blockIdx.x * CLT256_B + tid;
    char *smb = (char *)smem4;
    {
        uint32_t *sw = (uint32_t *)smem4;
        for (uint32_t p = 0; p < 256; p++) {
            const uint32_t v = planes[(size_t)CLT256_PERM[p] * T + t];
            const uint32_t G = p < 128 ? (p >> 2) : 64 + ((p - 128) >> 2);
            sw[(G * CLT256_B + tid) * 4 + (p & 3)] = v;
        }
    }
    __syncthreads();
    uint4 *lo_old = smem4 + tid;
    uint4 *lo_new = smem4 + 32 * CLT256_B + tid;
    uint4 *const hi = smem4 + 64 * CLT256_B + tid;
    uint4 *const stage = (uint4 *)(smb + CLT256_STATE_BYTES) + tid;
    // stream j of this thread starts at sample offset j*L from outp
    int8_t *const outp = out + (size_t)t * 32u * L * (OUT16 ? 2u : 1u);

#pragma unroll 1
    for (uint32_t it = 0; it < L; it++) {
#define CLT256S8_SECTION_BODY
#include "clt256_s8_gen.inc"
#undef CLT256S8_SECTION_BODY
        uint4 *tmp = lo_old;
        lo_old = lo_new;
        lo_new = tmp;
        if ((it & 7u) == 7u) {
            uint4 Z[8][2];
#pragma unroll
            for (int tau = 0; tau < 8; tau++) {
                Z[tau][0] = stage[(tau * 2) * CLT256_B];
                Z[tau][1] = stage[(tau * 2 + 1) * CLT256_B];
            }
            int8_t *const o8 = outp + (size_t)(it - 7u) * (OUT16 ? 2u : 1u);
#pragma unroll
            for (int s = 0; s < 8; s++) {
                uint32_t w[8], lo[4], hi4[4];
#pragma unroll
                for (int tau = 0; tau < 8; tau++) {
                    const uint4 q = Z[tau][s >> 2];
                    w[tau] = (s & 3) == 0 ? q.x : (s & 3) == 1 ? q.y : (s & 3) == 2 ? q.z : q.w;
                }
                bytes4x4(w[0], w[1], w[2], w[3], lo);
                bytes4x4(w[4], w[5], w[6], w[7], hi4);
#pragma unroll
                for (int m = 0; m < 4; m++) {
                    if (OUT16) {
                        const uint2 a = scale4(lo[m], noise_var), b = scale4(hi4[m], noise_var);
                        *(uint4 *)(o8 + (size_t)(8 * m + s) * L * 2u) = make_uint4(a.x, a.y, b.x, b.y);
                    } else {
                        *(uint2 *)(o8 + (size_t)(8 * m + s) * L) = make_uint2(lo[m], hi4[m]);
                    }
                }
            }
        }
    }
}

// One-pass variant of the sample-stream kernel (tools/gen_clt256_v2.py, kind "s8"): same clock as
// clt256_hist2_kernel; the tail of clock it-1 (popcount -> int8 -> 8x8 un-slice) runs inside iteration it.
// A stream's bytes are 1 per clock, and 1.2 M streams are open at a time: 8-byte pieces written
// straight to their final place (the two-pass kernel) leave partial sectors that fall out of L2 and
// cost a DRAM read-modify-write each (measured: 215 GB/s however fast the ALU side is).  Here the 8
// un-sliced words of a clock go to a global slab [clock % 32][word][thread] with coalesced 128-byte
// stores (it stays in L2: 1 KB per thread), and every 32 clocks a thread reads its slab column back,
// finishes the byte transpose and writes one full 32-byte sector per stream (64 B for int16).
#define CLT256V2_S8_K 32
#define CLT256V2_S8_STAGE(z0, z1, z2, z3, z4, z5, z6, z7)                                  \
    {                                                                                      \
        uint32_t *const sp = slab + (size_t)(((it - 1u) & (CLT256V2_S8_K - 1u)) * 8u) * T; \
        __stcg(sp, z0);                                                                    \
        __stcg(sp + (size_t)T, z1);                                                        \
        __stcg(sp + (size_t)2 * T, z2);                                                    \
        __stcg(sp + (size_t)3 * T, z3);                                                    \
        __stcg(sp + (size_t)4 * T, z4);                                                    \
        __stcg(sp + (size_t)5 * T, z5);                                                    \
        __stcg(sp + (size_t)6 * T, z6);                                                    \
        __stcg(sp + (size_t)7 * T, z7);                                                    \
    }

// clocks c0 .. c0+31 of this thread's 32 streams: slab -> final place
__device__ __forceinline__ void clt256v2_s8_flush(const uint32_t *slab, uint32_t T, int8_t *outp, uint32_t c0, uint32_t L,
                                                  bool OUT16, int32_t noise_var)
{
    int8_t *const o8 = outp + (size_t)c0 * (OUT16 ? 2u : 1u);
#pragma unroll 1
    for (int s = 0; s < 8; s++) {
        uint32_t W[CLT256V2_S8_K], o[CLT256V2_S8_K / 4][4];
#pragma unroll
        for (int tau = 0; tau < CLT256V2_S8_K; tau++) W[tau] = __ldcg(slab + (size_t)(tau * 8 + s) * T);
#pragma unroll
        for (int g = 0; g < CLT256V2_S8_K / 4; g++) bytes4x4(W[4 * g], W[4 * g + 1], W[4 * g + 2], W[4 * g + 3], o[g]);
#pragma unroll
        for (int m = 0; m < 4; m++) {          // byte lane m of word s = stream 8m+s
            if (OUT16) {
                uint4 *const d = (uint4 *)(o8 + (size_t)(8 * m + s) * L * 2u);
#pragma unroll
                for (int g = 0; g < CLT256V2_S8_K / 4; g += 2) {
                    const uint2 a = scale4(o[g][m], noise_var), b = scale4(o[g + 1][m], noise_var);
                    __stcs(d + g / 2, make_uint4(a.x, a.y, b.x, b.y));
                }
            } else {
                uint4 *const d = (uint4 *)(o8 + (size_t)(8 * m + s) * L);
#pragma unroll
                for (int g = 0; g < CLT256V2_S8_K / 4; g += 4)
                    __stcs(d + g / 4, make_uint4(o[g][m], o[g + 1][m], o[g + 2][m], o[g + 3][m]));
            }
        }
    }
}

__global__ void __launch_bounds__(CLT256V2_B, CLT256V2_CTAS_PER_SM)
clt256_s8v2_kernel(const uint32_t *__restrict__ planes, uint32_t T, uint32_t L, int8_t *__restrict__ out,
                   int32_t noise_var, uint32_t *__restrict__ slab_base)
{
    const bool OUT16 = noise_var >= 0;
    extern __shared__ uint4 smem4[];
    const uint32_t tid = threadIdx.x;
    const uint32_t t = blockIdx.x * CLT256V2_B + tid;
    uint32_t *const st = (uint32_t *)smem4 + tid;
    for (uint32_t p = 0; p < CLT256V2_NSLOTS; p++) st[p * CLT256V2_B] = planes[(size_t)CLT256V2_SLOT_BIT[p] * T + t];
    uint32_t *const slab = slab_base + t;
    int8_t *const outp = out + (size_t)t * 32u * L * (OUT16 ? 2u : 1u);   // stream j: sample offset j*L
#define CLT256V2_SECTION_RESIDENT_DECL
#include "clt256_v2_gen.inc"
#undef CLT256V2_SECTION_RESIDENT_DECL

#pragma unroll 1
    for (uint32_t it = 0; it < L; it++) {
#define CLT256V2_SECTION_S8_BODY
#include "clt256_v2_gen.inc"
#undef CLT256V2_SECTION_S8_BODY
        if ((it & (CLT256V2_S8_K - 1u)) == 0u && it != 0u)
            clt256v2_s8_flush(slab, T, outp, it - CLT256V2_S8_K, L, OUT16, noise_var);
    }
    {
        const uint32_t it = L;       // the last clock's tail (L is a multiple of 32: it lands in slab row 31)
#define CLT256V2_SECTION_S8_TAIL
#include "clt256_v2_gen.inc"
#undef CLT256V2_SECTION_S8_TAIL
        clt256v2_s8_flush(slab, T, outp, L - CLT256V2_S8_K, L, OUT16, noise_var);
    }
}

// Decimated sample stream: out[j] = g[j * spb + delay] of every stream -- what the receiver keeps of the
// noise (rx.py:18-19 takes sample `delay` of the spb samples of a bit).  The recurrence has to be clocked
// through all spb samples, but only the sampled clock needs the CLT tree: that clock is the "s8" body
// (popcount network, un-slice, slab), the other spb-1 are the clock alone ("plain" body: 492 LOP3, the
// loads and stores of the non-resident planes).  Ld = decimated samples per stream (a multiple of 32).
// CLT256_DEC_P2 (the default; 0 = the round-1 kernel): the plain clocks keep 151 planes in registers instead of 104 -- the popcount network's registers
// are free there -- and the 47 extra ones are spilled to / filled from their shared-memory slots around the sampled
// clock.  ncu on the 104-plane version: LSU data pipe 82 % busy, ALU pipe 74 %: the 152 LDS + 152 STS of every plain
// clock were the limit.  (A compile-time switch and not a template parameter: cudafe++ of CUDA 12.9 does not finish on
// the templated version of this kernel.)
#ifndef CLT256_DEC_P2
#define CLT256_DEC_P2 1
#endif
__global__ void __launch_bounds__(CLT256V2_B, CLT256V2_CTAS_PER_SM)
clt256_dec_kernel(const uint32_t *__restrict__ planes, uint32_t T, uint32_t Ld, uint32_t spb, uint32_t delay,
                  int8_t *__restrict__ out, uint32_t *__restrict__ slab_base)
{
    extern __shared__ uint4 smem4[];
    const uint32_t tid = threadIdx.x;
    const uint32_t t = blockIdx.x * CLT256V2_B + tid;
    uint32_t *const st = (uint32_t *)smem4 + tid;
    for (uint32_t p = 0; p < CLT256V2_NSLOTS; p++) st[p * CLT256V2_B] = planes[(size_t)CLT256V2_SLOT_BIT[p] * T + t];
    uint32_t *const slab = slab_base + t;
    int8_t *const outp = out + (size_t)t * 32u * Ld;          // stream j: decimated sample offset j*Ld
#define CLT256V2_SECTION_RESIDENT_DECL
#include "clt256_v2_gen.inc"
#undef CLT256V2_SECTION_RESIDENT_DECL

#if CLT256_DEC_P2
#define CLT256V2_SECTION_P2_DECL
#include "clt256_v2_gen.inc"
#undef CLT256V2_SECTION_P2_DECL
#endif

#pragma unroll 1
    for (uint32_t it = 0; it < Ld; it++) {
#pragma unroll 1
        for (uint32_t ph = 0; ph < spb; ph++) {
            if (ph == delay) {
#if CLT256_DEC_P2
#define CLT256V2_SECTION_P2_SPILL
#include "clt256_v2_gen.inc"
#undef CLT256V2_SECTION_P2_SPILL
#endif
#define CLT256V2_SECTION_S8_BODY
#include "clt256_v2_gen.inc"
#undef CLT256V2_SECTION_S8_BODY
#if CLT256_DEC_P2
#define CLT256V2_SECTION_P2_FILL
#include "clt256_v2_gen.inc"
#undef CLT256V2_SECTION_P2_FILL
#endif
            } else {
#if CLT256_DEC_P2
#define CLT256V2_SECTION_P2_BODY
#include "clt256_v2_gen.inc"
#undef CLT256V2_SECTION_P2_BODY
#else
#define CLT256V2_SECTION_P_BODY
#include "clt256_v2_gen.inc"
#undef CLT256V2_SECTION_P_BODY
#endif
            }
        }
        if ((it & (CLT256V2_S8_K - 1u)) == 0u && it != 0u)
            clt256v2_s8_flush(slab, T, outp, it - CLT256V2_S8_K, Ld, false, 0);
    }
    {
        const uint32_t it = Ld;      // the last sampled clock's tail (Ld is a multiple of 32: slab row 31)
#define CLT256V2_SECTION_S8_TAIL
#include "clt256_v2_gen.inc"
#undef CLT256V2_SECTION_S8_TAIL
        clt256v2_s8_flush(slab, T, outp, Ld - CLT256V2_S8_K, Ld, false, 0);
    }
}

int bbb_clt_spec_dec_ok(const bbb_ctx *c)
{
    return c->shipped == 256 && !c->tune_no_spec && c->tune_spec_variant != 1;
}

int bbb_clt_spec_dec_launch(bbb_ctx *c, uint32_t blocks, uint32_t Ld, uint32_t spb, uint32_t delay, const uint32_t *planes,
                            int8_t *out)
{
    const uint32_t T = blocks * CLT256V2_B;
    uint32_t *slab = (uint32_t *)c->scratch(11, (size_t)CLT256V2_S8_K * 8 * T * 4);
    if (!slab) return BBB_ECUDA;
    auto kern = clt256_dec_kernel;
    BBB_CUDA(c, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, CLT256V2_STATE_BYTES));
    kern<<<blocks, CLT256V2_B, CLT256V2_STATE_BYTES, c->stream>>>(planes, T, Ld, spb, delay, out, slab);
    c->last_kernel = "clt256_dec";
    c->last_launches++;
    BBB_CUDA(c, cudaGetLastError());
    return BBB_OK;
}

// raw URNG words (LUTOPT.x after every clock, rng.py:27): 32 x 32 bit transposes of the new planes,
// four 32-bit words of a stream leave as one 128-bit store (a stream's clock = one 32-byte sector)
#define CLT256W_SECTION_PERM
#include "clt256_words_gen.inc"
#undef CLT256W_SECTION_PERM

#define CLT256_W_EMIT(w, r0, r1, r2, r3, r4, r5, r6, r7, r8, r9, r10, r11, r12, r13, r14, r15, r16, r17, r18, r19, \
                      r20, r21, r22, r23, r24, r25, r26, r27, r28, r29, r30, r31)                                  \
    {                                                                                                              \
        uint32_t tw[32] = {r0, r1, r2, r3, r4, r5, r6, r7, r8, r9, r10, r11, r12, r13, r14, r15, r16, r17, r18,     \
                           r19, r20, r21, r22, r23, r24, r25, r26, r27, r28, r29, r30, r31};                        \
        transpose32_fast(tw);                                                                                      \
        _Pragma("unroll") for (int j = 0; j < 32; j++) wb[(w) & 3][j] = tw[j];                                     \
    }

#define CLT256_W_STORE(which)                                                                          \
    {                                                                                                  \
        uint32_t *const o32 = outp + (size_t)it * 8u + (which) * 4u;                                   \
        _Pragma("unroll") for (int j = 0; j < 32; j++)                                                 \
            __stcs((uint4 *)(o32 + (size_t)j * L * 8u), make_uint4(wb[0][j], wb[1][j], wb[2][j], wb[3][j])); \
    }

__global__ void __launch_bounds__(CLT256_B, 1)
clt256_words_kernel(const uint32_t *__restrict__ planes, uint32_t T, uint32_t L, uint32_t *__restrict__ out)
{
    extern __shared__ uint4 smem4[];
    const uint32_t tid = threadIdx.x;
    const uint32_t t = blockIdx.x * CLT256_B + tid;
    {
        uint32_t *sw = (uint32_t *)smem4;
        for (uint32_t p = 0; p < 256; p++) {
            const uint32_t v = planes[(size_t)CLT256W_PERM[p] * T + t];
            const uint32_t G = p < 128 ? (p >> 2) : 64 + ((p - 128) >> 2);
            sw[(G * CLT256_B + tid) * 4 + (p & 3)] = v;
        }
    }
    __syncthreads();
    uint4 *lo_old = smem4 + tid;
    uint4 *lo_new = smem4 + 32 * CLT256_B + tid;
    uint4 *const hi = smem4 + 64 * CLT256_B + tid;
    uint32_t *const outp = out + (size_t)t * 32u * L * 8u;   // stream j, clock it: outp + (j*L + it)*8
    uint32_t wb[4][32];

#pragma unroll 1
    for (uint32_t it = 0; it < L; it++) {
#define CLT256W_SECTION_BODY
#include "clt256_words_gen.inc"
#undef CLT256W_SECTION_BODY
        uint4 *tmp = lo_old;
        lo_old = lo_new;
        lo_new = tmp;
    }
}

// One-pass variant of the raw-word kernel (tools/gen_clt256_v2.py, kind "words"): the one-pass clock without
// the popcount network; then the NEW state is un-sliced 32 planes at a time (32x32 register transposes,
// two PRMT stages + three shift/LOP3 stages) into a thread-private shared-memory stage [stream][word],
// and every stream's 8 words leave as one complete 32-byte sector (two adjacent 128-bit stores).  The
// two-pass kernel stored the two halves of a sector from different passes: 606 k half-written sectors
// in flight, each costing a DRAM read-modify-write (ncu: 28 % ALU, 3 long-scoreboard stalls per issue).
// One CTA per SM: state 76 KB + stage 128 KB.
#define CLT256V2W_STAGE_WORDS (32 * 8 * CLT256V2_B)
#define CLT256V2W_SMEM (CLT256V2_STATE_BYTES + CLT256V2W_STAGE_WORDS * 4)

// stage word (stream j, word w) of thread `tid` lives in column tid ^ 16*(w >> 2): the two lanes that later
// fetch the two halves of one thread's 32-byte piece then hit different banks.
// The eight emits of a clock run as ONE loop: a switch copies the 32 source planes of word w into tw[], the
// transpose and the staging stores exist once.  (Straight-line, the loop body of this kernel was ~5500 instructions
// = 88 KB: ncu showed 0.57 no-instruction stalls per issued instruction, four warps streaming different parts of it.)
#define CLT256V2_W_BEGIN                                                                                             \
    __syncwarp();                        /* every lane has drained the previous clock's words */                     \
    _Pragma("unroll 1") for (uint32_t w_ = 0; w_ < 8u; w_++) {                                                       \
        uint32_t tw[32];                                                                                             \
        switch (w_) {
#define CLT256V2_W_EMIT(w, r0, r1, r2, r3, r4, r5, r6, r7, r8, r9, r10, r11, r12, r13, r14, r15, r16, r17, r18, r19, \
                        r20, r21, r22, r23, r24, r25, r26, r27, r28, r29, r30, r31)                                  \
        case w: {                                                                                                    \
            const uint32_t src[32] = {r0, r1, r2, r3, r4, r5, r6, r7, r8, r9, r10, r11, r12, r13, r14, r15, r16,     \
                                      r17, r18, r19, r20, r21, r22, r23, r24, r25, r26, r27, r28, r29, r30, r31};    \
            _Pragma("unroll") for (int i = 0; i < 32; i++) tw[i] = src[i];                                           \
        } break;
#define CLT256V2_W_END                                                                                               \
        }                                                                                                            \
        transpose32_fast(tw);                                                                                        \
        const uint32_t sa = (w_ < 4u ? stage_a : stage_b) + w_ * (uint32_t)(CLT256V2_B * 4);                         \
        _Pragma("unroll") for (int j = 0; j < 32; j++)                                                               \
            asm volatile("st.shared.u32 [%0], %1;" ::"r"(sa + (uint32_t)(j * 8 * CLT256V2_B * 4)), "r"(tw[j]));      \
    }

// One stream index j of the PREVIOUS clock leaves the stage (the generator places 32 of these between the steps of
// the clock, so the stores drain behind the clock's ALU work instead of stalling the warp after the transposes:
// ncu on the unpipelined kernel showed 1.25 long-scoreboard stalls per issued instruction and 26 % issue activity).
#define CLT256V2_W_DRAIN(j)                                                                                          \
    if (it != 0u) {                                                                                                  \
        uint32_t v[8];                                                                                               \
        _Pragma("unroll") for (int i = 0; i < 4; i++) {                                                              \
            asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v[i]) : "r"(rd0 + (uint32_t)(((j) * 8 + i) * CLT256V2_B * 4)));     \
            asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v[4 + i]) : "r"(rd1 + (uint32_t)(((j) * 8 + i) * CLT256V2_B * 4))); \
        }                                                                                                            \
        const size_t off = ((size_t)(j) * L + (it - 1u)) * 8u;                                                       \
        CLT256V2W_STORE((uint4 *)(op0 + off), make_uint4(v[0], v[1], v[2], v[3]));                                   \
        CLT256V2W_STORE((uint4 *)(op1 + off), make_uint4(v[4], v[5], v[6], v[7]));                                   \
    }

// Store policy of the final 16-byte pieces.  A stream's 128-byte line is completed by four consecutive clocks
// (one 32-byte sector each), and 606 k streams are open at once: with evict-first (streaming) stores the sectors
// reach HBM one by one, scattered over as many DRAM pages; keeping the lines in L2 (evict_last, 77 MB in flight)
// lets them leave as full lines.  Measured at 2^28 steps: 0 (__stcs) 1350 GB/s, 1 (default) 1350, 2 (evict_last) 1662.
#ifndef CLT256V2W_STORE_KIND
#define CLT256V2W_STORE_KIND 2
#endif
#if CLT256V2W_STORE_KIND == 0
#define CLT256V2W_STORE(p, v) __stcs(p, v)
#elif CLT256V2W_STORE_KIND == 1
#define CLT256V2W_STORE(p, v) (*(p) = (v))
#else
__device__ __forceinline__ void st_evict_last(uint4 *p, uint4 v)
{
    uint64_t pol;
    asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(pol));
    asm volatile("st.global.L2::cache_hint.v4.u32 [%0], {%1, %2, %3, %4}, %5;" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w),
                 "l"(pol) : "memory");
}
#define CLT256V2W_STORE(p, v) st_evict_last(p, v)
#endif
__device__ __forceinline__ void st_policy(uint4 *p, uint4 v, uint64_t pol)
{
    asm volatile("st.global.L2::cache_hint.v4.u32 [%0], {%1, %2, %3, %4}, %5;" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w),
                 "l"(pol) : "memory");
}

__global__ void __launch_bounds__(CLT256V2_B, 1)
clt256_words2_kernel(const uint32_t *__restrict__ planes, uint32_t T, uint32_t L, uint32_t *__restrict__ out)
{
    extern __shared__ uint4 smem4[];
    const uint32_t tid = threadIdx.x;
    const uint32_t t = blockIdx.x * CLT256V2_B + tid;
    uint32_t *const st = (uint32_t *)smem4 + tid;
    for (uint32_t p = 0; p < CLT256V2_NSLOTS; p++) st[p * CLT256V2_B] = planes[(size_t)CLT256V2_SLOT_BIT[p] * T + t];
    const uint32_t stage_base = (uint32_t)__cvta_generic_to_shared((char *)smem4 + CLT256V2_STATE_BYTES);
    const uint32_t stage_a = stage_base + tid * 4u, stage_b = stage_base + (tid ^ 16u) * 4u;
    // store side: lanes 2q, 2q+1 write the two 16-byte halves of ONE stream's 32-byte clock, so a warp store covers
    // 16 complete sectors (it was 32 half sectors when every lane stored its own thread's streams): lane (q, h)
    // serves threads wbase + q and wbase + q + 16 of its warp
    const uint32_t lane = tid & 31u, q = lane >> 1, h = lane & 1u, wbase = tid & ~31u;
    const uint32_t u0 = wbase + q, u1 = u0 + 16u;
    const uint32_t rd0 = stage_base + ((4u * h) * CLT256V2_B + (u0 ^ (16u * h))) * 4u;
    const uint32_t rd1 = stage_base + ((4u * h) * CLT256V2_B + (u1 ^ (16u * h))) * 4u;
    uint32_t *const op0 = out + (size_t)(blockIdx.x * CLT256V2_B + u0) * 32u * L * 8u + 4u * h;   // stream j, clock it: + (j*L + it)*8
    uint32_t *const op1 = out + (size_t)(blockIdx.x * CLT256V2_B + u1) * 32u * L * 8u + 4u * h;
#define CLT256V2_SECTION_RESIDENT_DECL
#include "clt256_v2_gen.inc"
#undef CLT256V2_SECTION_RESIDENT_DECL

#pragma unroll 1
    for (uint32_t it = 0; it < L; it++) {
        __syncwarp();                    // the stage holds clock it-1 of every lane's threads
#define CLT256V2_SECTION_W_BODY
#include "clt256_v2_gen.inc"
#undef CLT256V2_SECTION_W_BODY
    }
    __syncwarp();
    {                                    // the last clock's words
        const uint32_t it = L;
#pragma unroll 4
        for (int j = 0; j < 32; j++) { CLT256V2_W_DRAIN(j) }
    }
}

// ---------------------------------------------------------------------------
// clt256_words4_kernel: the raw-word stream with TWO CTAs per SM.  words2 keeps a whole clock of un-sliced words
// (32 streams x 8 words x 128 threads = 128 KB) in shared memory, so only four warps fit an SM and each scheduler has
// one warp whose dependency and shared-memory latencies nobody hides (ncu: issue slots 31 % used, ALU 37 %).  Here a
// clock is un-sliced in four groups of 8 streams: for group g and word w the kernel takes byte g of the 32 planes of
// the word (24 PRMT pack them into 8 words, byte b of word k = plane 8b+k) and transposes the 8x8 bit matrices
// (3 stages x 4 pairs), which yields word w of the 8 streams directly -- 72 ALU instructions where a full 32x32
// transpose costs 256 for 32 streams, so the total is the same.  The stage holds 8 streams x 8 words (32 KB), lane
// pairs then store complete 32-byte sectors as in words2, and state + stage = 108 KB: two CTAs, eight warps per SM.
// ---------------------------------------------------------------------------
#pragma push_macro("CLT256V2_W_DRAIN")
#pragma push_macro("CLT256V2_W_BEGIN")
#pragma push_macro("CLT256V2_W_EMIT")
#pragma push_macro("CLT256V2_W_END")
#undef CLT256V2_W_DRAIN
#undef CLT256V2_W_BEGIN
#undef CLT256V2_W_EMIT
#undef CLT256V2_W_END
#define CLT256V4W_STAGE_BYTES (8 * 8 * CLT256V2_B * 4)
#define CLT256V4W_SMEM (CLT256V2_STATE_BYTES + CLT256V4W_STAGE_BYTES)

__device__ __forceinline__ uint32_t prmt_r(uint32_t a, uint32_t b, uint32_t sel)
{
    uint32_t d;
    asm("prmt.b32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(sel));
    return d;
}

// tw[i] = plane i of a word (bit j = stream j).  Returns in o[s] the word of stream 8g+s: bit i = bit 8g+s of tw[i].
__device__ __forceinline__ void unslice8(const uint32_t (&tw)[32], uint32_t sel2, uint32_t (&o)[8])
{
#pragma unroll
    for (int k = 0; k < 8; k++) {
        const uint32_t lo = prmt_r(tw[k], tw[8 + k], sel2), hi = prmt_r(tw[16 + k], tw[24 + k], sel2);
        o[k] = __byte_perm(lo, hi, 0x5410);          // byte b = byte g of plane 8b + k
    }
#pragma unroll
    for (int st = 0; st < 3; st++) {
        const int j = 4 >> st;
        const uint32_t m = st == 0 ? 0x0F0F0F0Fu : st == 1 ? 0x33333333u : 0x55555555u;
#pragma unroll
        for (int k = 0; k < 8; k++) {
            if (k & j) continue;
            const uint32_t x = o[k], y = o[k + j];
            o[k] = (x & m) | ((y << j) & ~m);
            o[k + j] = ((x >> j) & m) | (y & ~m);
        }
    }
}

#define CLT256V2_W_DRAIN(j)
#define CLT256V2_W_BEGIN                                                                                             \
    _Pragma("unroll 1") for (uint32_t g_ = 0; g_ < 4u; g_++) {                                                       \
        const uint32_t sel2 = g_ | ((4u + g_) << 4);                                                                 \
        _Pragma("unroll 1") for (uint32_t w_ = 0; w_ < 8u; w_++) {                                                   \
            uint32_t tw[32];                                                                                         \
            switch (w_) {
#define CLT256V2_W_EMIT(w, r0, r1, r2, r3, r4, r5, r6, r7, r8, r9, r10, r11, r12, r13, r14, r15, r16, r17, r18, r19, \
                        r20, r21, r22, r23, r24, r25, r26, r27, r28, r29, r30, r31)                                  \
            case w: {                                                                                                \
                const uint32_t src[32] = {r0, r1, r2, r3, r4, r5, r6, r7, r8, r9, r10, r11, r12, r13, r14, r15, r16, \
                                          r17, r18, r19, r20, r21, r22, r23, r24, r25, r26, r27, r28, r29, r30, r31}; \
                _Pragma("unroll") for (int i = 0; i < 32; i++) tw[i] = src[i];                                       \
            } break;
#define CLT256V2_W_END                                                                                               \
            }                                                                                                        \
            uint32_t o[8];                                                                                           \
            unslice8(tw, sel2, o);                                                                                   \
            const uint32_t sa = (w_ < 4u ? stage_a : stage_b) + w_ * (uint32_t)(CLT256V2_B * 4);                     \
            _Pragma("unroll") for (int s8 = 0; s8 < 8; s8++)                                                         \
                asm volatile("st.shared.u32 [%0], %1;" ::"r"(sa + (uint32_t)(s8 * 8 * CLT256V2_B * 4)), "r"(o[s8])); \
        }                                                                                                            \
        __syncwarp();                                                                                                \
        _Pragma("unroll") for (int s8 = 0; s8 < 8; s8++) {                                                           \
            uint32_t v[8];                                                                                           \
            _Pragma("unroll") for (int i = 0; i < 4; i++) {                                                          \
                asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v[i]) : "r"(rd0 + (uint32_t)((s8 * 8 + i) * CLT256V2_B * 4)));     \
                asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v[4 + i]) : "r"(rd1 + (uint32_t)((s8 * 8 + i) * CLT256V2_B * 4))); \
            }                                                                                                        \
            const size_t off = ((size_t)(g_ * 8u + s8) * L + it) * 8u;                                               \
            CLT256V2W_STORE((uint4 *)(op0 + off), make_uint4(v[0], v[1], v[2], v[3]));                               \
            CLT256V2W_STORE((uint4 *)(op1 + off), make_uint4(v[4], v[5], v[6], v[7]));                               \
        }                                                                                                            \
        __syncwarp();                                                                                                \
    }

__global__ void __launch_bounds__(CLT256V2_B, 2)
clt256_words4_kernel(const uint32_t *__restrict__ planes, uint32_t T, uint32_t L, uint32_t *__restrict__ out)
{
    extern __shared__ uint4 smem4[];
    const uint32_t tid = threadIdx.x;
    const uint32_t t = blockIdx.x * CLT256V2_B + tid;
    uint32_t *const st = (uint32_t *)smem4 + tid;
    for (uint32_t p = 0; p < CLT256V2_NSLOTS; p++) st[p * CLT256V2_B] = planes[(size_t)CLT256V2_SLOT_BIT[p] * T + t];
    const uint32_t stage_base = (uint32_t)__cvta_generic_to_shared((char *)smem4 + CLT256V2_STATE_BYTES);
    const uint32_t stage_a = stage_base + tid * 4u, stage_b = stage_base + (tid ^ 16u) * 4u;
    const uint32_t lane = tid & 31u, q = lane >> 1, h = lane & 1u, wbase = tid & ~31u;
    const uint32_t u0 = wbase + q, u1 = u0 + 16u;
    const uint32_t rd0 = stage_base + ((4u * h) * CLT256V2_B + (u0 ^ (16u * h))) * 4u;
    const uint32_t rd1 = stage_base + ((4u * h) * CLT256V2_B + (u1 ^ (16u * h))) * 4u;
    uint32_t *const op0 = out + (size_t)(blockIdx.x * CLT256V2_B + u0) * 32u * L * 8u + 4u * h;   // stream j, clock it: + (j*L + it)*8
    uint32_t *const op1 = out + (size_t)(blockIdx.x * CLT256V2_B + u1) * 32u * L * 8u + 4u * h;
#define CLT256V2_SECTION_RESIDENT_DECL
#include "clt256_v2_gen.inc"
#undef CLT256V2_SECTION_RESIDENT_DECL

#pragma unroll 1
    for (uint32_t it = 0; it < L; it++) {
#define CLT256V2_SECTION_W_BODY
#include "clt256_v2_gen.inc"
#undef CLT256V2_SECTION_W_BODY
    }
}
#pragma pop_macro("CLT256V2_W_END")
#pragma pop_macro("CLT256V2_W_EMIT")
#pragma pop_macro("CLT256V2_W_BEGIN")
#pragma pop_macro("CLT256V2_W_DRAIN")

// ---------------------------------------------------------------------------
// clt256_words5_kernel (knob words_variant = 5): words2 with the un-sliced clock staged in an L2-resident global
// slab [word][stream][thread] (coalesced 128-byte rows in both directions: 1 KB per thread, 39 MB for the grid)
// instead of 128 KB of shared memory, which leaves room for two CTAs = eight warps per SM.  The lane pairs read
// their halves back with L1-bypassing loads and store complete sectors exactly as in words2.
// ---------------------------------------------------------------------------
#pragma push_macro("CLT256V2_W_DRAIN")
#pragma push_macro("CLT256V2_W_BEGIN")
#pragma push_macro("CLT256V2_W_EMIT")
#pragma push_macro("CLT256V2_W_END")
#undef CLT256V2_W_DRAIN
#undef CLT256V2_W_BEGIN
#undef CLT256V2_W_EMIT
#undef CLT256V2_W_END
#define CLT256V2_W_DRAIN(j)                                                                                          \
    if (it != 0u) {                                                                                                  \
        uint32_t v[8];                                                                                               \
        _Pragma("unroll") for (int i = 0; i < 4; i++) {                                                              \
            v[i] = __ldcg(rd0 + (size_t)((4u * h + i) * 32u + (j)) * T);                                             \
            v[4 + i] = __ldcg(rd1 + (size_t)((4u * h + i) * 32u + (j)) * T);                                         \
        }                                                                                                            \
        const size_t off = ((size_t)(j) * L + (it - 1u)) * 8u;                                                       \
        CLT256V2W_STORE((uint4 *)(op0 + off), make_uint4(v[0], v[1], v[2], v[3]));                                   \
        CLT256V2W_STORE((uint4 *)(op1 + off), make_uint4(v[4], v[5], v[6], v[7]));                                   \
    }
#define CLT256V2_W_BEGIN                                                                                             \
    __syncwarp();                        /* every lane has drained the previous clock's words */                     \
    _Pragma("unroll 1") for (uint32_t w_ = 0; w_ < 8u; w_++) {                                                       \
        uint32_t tw[32];                                                                                             \
        switch (w_) {
#define CLT256V2_W_EMIT(w, r0, r1, r2, r3, r4, r5, r6, r7, r8, r9, r10, r11, r12, r13, r14, r15, r16, r17, r18, r19, \
                        r20, r21, r22, r23, r24, r25, r26, r27, r28, r29, r30, r31)                                  \
        case w: {                                                                                                    \
            const uint32_t src[32] = {r0, r1, r2, r3, r4, r5, r6, r7, r8, r9, r10, r11, r12, r13, r14, r15, r16,     \
                                      r17, r18, r19, r20, r21, r22, r23, r24, r25, r26, r27, r28, r29, r30, r31};    \
            _Pragma("unroll") for (int i = 0; i < 32; i++) tw[i] = src[i];                                           \
        } break;
#define CLT256V2_W_END                                                                                               \
        }                                                                                                            \
        transpose32_fast(tw);                                                                                        \
        uint32_t *const sw = slab_w + (size_t)w_ * 32u * T;                                                          \
        _Pragma("unroll") for (int j = 0; j < 32; j++) __stcg(sw + (size_t)j * T, tw[j]);                            \
    }

__global__ void __launch_bounds__(CLT256V2_B, 2)
clt256_words5_kernel(const uint32_t *__restrict__ planes, uint32_t T, uint32_t L, uint32_t *__restrict__ out,
                     uint32_t *__restrict__ slab)
{
    extern __shared__ uint4 smem4[];
    const uint32_t tid = threadIdx.x;
    const uint32_t t = blockIdx.x * CLT256V2_B + tid;
    uint32_t *const st = (uint32_t *)smem4 + tid;
    for (uint32_t p = 0; p < CLT256V2_NSLOTS; p++) st[p * CLT256V2_B] = planes[(size_t)CLT256V2_SLOT_BIT[p] * T + t];
    const uint32_t lane = tid & 31u, q = lane >> 1, h = lane & 1u, wbase = tid & ~31u;
    const uint32_t u0 = wbase + q, u1 = u0 + 16u;
    uint32_t *const slab_w = slab + t;                                        // this thread's column of the slab
    const uint32_t *const rd0 = slab + blockIdx.x * CLT256V2_B + u0;          // the columns this lane stores for
    const uint32_t *const rd1 = slab + blockIdx.x * CLT256V2_B + u1;
    uint32_t *const op0 = out + (size_t)(blockIdx.x * CLT256V2_B + u0) * 32u * L * 8u + 4u * h;   // stream j, clock it: + (j*L + it)*8
    uint32_t *const op1 = out + (size_t)(blockIdx.x * CLT256V2_B + u1) * 32u * L * 8u + 4u * h;
#define CLT256V2_SECTION_RESIDENT_DECL
#include "clt256_v2_gen.inc"
#undef CLT256V2_SECTION_RESIDENT_DECL

#pragma unroll 1
    for (uint32_t it = 0; it < L; it++) {
        __syncwarp();                    // the slab holds clock it-1 of every lane's threads
#define CLT256V2_SECTION_W_BODY
#include "clt256_v2_gen.inc"
#undef CLT256V2_SECTION_W_BODY
    }
    __syncwarp();
    {
        const uint32_t it = L;
#pragma unroll 4
        for (int j = 0; j < 32; j++) { CLT256V2_W_DRAIN(j) }
    }
}
#pragma pop_macro("CLT256V2_W_END")
#pragma pop_macro("CLT256V2_W_EMIT")
#pragma pop_macro("CLT256V2_W_BEGIN")
#pragma pop_macro("CLT256V2_W_DRAIN")

// ---------------------------------------------------------------------------
// clt256_words6_kernel (knob words_variant = 6): words2 with the drain moved to its own warps.  The source page of
// the words2 capture (profiles/r2_words2_ncu_full.txt) puts 59 % of its stall samples in the recurrence part of the
// clock although that part holds only 2064 of the 5260 instructions: the 32 drain pieces woven into it (8 LDS ->
// 2 STG each) expose the shared-memory latency at every store, and the next piece's loads wait for the scattered
// stores to read their registers (long-scoreboard stalls on LOP3 / LDS), with one warp per scheduler and nobody to
// hide either.  Here warps 0-3 run the clock and the transposes exactly as in words2, and warps 4-7 (255 registers
// each as well: 256 x 255 fits the register file) copy stage -> HBM, four streams per batch.  Producer / consumer
// hand-off per warp pair over two named barriers (full, empty), PTX bar.arrive / bar.sync with 64 participants; the
// stage stays single-buffered (128 KB), so the drain overlaps the NEXT clock's recurrence and the transposes wait
// for it.
// ---------------------------------------------------------------------------
#pragma push_macro("CLT256V2_W_DRAIN")
#pragma push_macro("CLT256V2_W_BEGIN")
#pragma push_macro("CLT256V2_W_END")
#undef CLT256V2_W_DRAIN
#undef CLT256V2_W_BEGIN
#undef CLT256V2_W_END
#define CLT256V2_W_DRAIN(j)
#define CLT256V2_W_BEGIN                                                                                             \
    if (it != 0u) asm volatile("bar.sync %0, 64;" ::"r"(bar_empty) : "memory");   /* clock it-1 has left the stage */ \
    _Pragma("unroll 1") for (uint32_t w_ = 0; w_ < 8u; w_++) {                                                       \
        uint32_t tw[32];                                                                                             \
        switch (w_) {
#define CLT256V2_W_END                                                                                               \
        }                                                                                                            \
        transpose32_fast(tw);                                                                                        \
        const uint32_t sa = (w_ < 4u ? stage_a : stage_b) + w_ * (uint32_t)(CLT256V2_B * 4);                         \
        _Pragma("unroll") for (int j = 0; j < 32; j++)                                                               \
            asm volatile("st.shared.u32 [%0], %1;" ::"r"(sa + (uint32_t)(j * 8 * CLT256V2_B * 4)), "r"(tw[j]));      \
    }                                                                                                                \
    asm volatile("bar.arrive %0, 64;" ::"r"(bar_full) : "memory");

__global__ void __launch_bounds__(2 * CLT256V2_B, 1)
clt256_words6_kernel(const uint32_t *__restrict__ planes, uint32_t T, uint32_t L, uint32_t *__restrict__ out)
{
    extern __shared__ uint4 smem4[];
    const uint32_t stage_base = (uint32_t)__cvta_generic_to_shared((char *)smem4 + CLT256V2_STATE_BYTES);
    const uint32_t pair = (threadIdx.x >> 5) & 3u;                   // compute warp `pair` feeds drain warp 4 + `pair`
    const uint32_t bar_full = 1u + pair, bar_empty = 5u + pair;      // barrier 0 is __syncthreads
    if (threadIdx.x >= CLT256V2_B) {
        // drain warp: lanes 2q, 2q+1 write the two 16-byte halves of one stream's clock (words2's store side)
        const uint32_t lane = threadIdx.x & 31u, q = lane >> 1, h = lane & 1u;
        const uint32_t u0 = pair * 32u + q, u1 = u0 + 16u;
        const uint32_t rd0 = stage_base + ((4u * h) * CLT256V2_B + (u0 ^ (16u * h))) * 4u;
        const uint32_t rd1 = stage_base + ((4u * h) * CLT256V2_B + (u1 ^ (16u * h))) * 4u;
        uint32_t *const op0 = out + (size_t)(blockIdx.x * CLT256V2_B + u0) * 32u * L * 8u + 4u * h;
        uint32_t *const op1 = out + (size_t)(blockIdx.x * CLT256V2_B + u1) * 32u * L * 8u + 4u * h;
#pragma unroll 1
        for (uint32_t it = 0; it < L; it++) {
            asm volatile("bar.sync %0, 64;" ::"r"(bar_full) : "memory");
#pragma unroll 2
            for (uint32_t jb = 0; jb < 32u; jb += 4u) {
                uint32_t v[4][8];
#pragma unroll
                for (uint32_t jj = 0; jj < 4u; jj++)
#pragma unroll
                    for (uint32_t i = 0; i < 4u; i++) {
                        const uint32_t o = ((jb + jj) * 8u + i) * (uint32_t)(CLT256V2_B * 4);
                        asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v[jj][i]) : "r"(rd0 + o));
                        asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v[jj][4 + i]) : "r"(rd1 + o));
                    }
#pragma unroll
                for (uint32_t jj = 0; jj < 4u; jj++) {
                    const size_t off = ((size_t)(jb + jj) * L + it) * 8u;
                    CLT256V2W_STORE((uint4 *)(op0 + off), make_uint4(v[jj][0], v[jj][1], v[jj][2], v[jj][3]));
                    CLT256V2W_STORE((uint4 *)(op1 + off), make_uint4(v[jj][4], v[jj][5], v[jj][6], v[jj][7]));
                }
            }
            if (it + 1u < L) asm volatile("bar.arrive %0, 64;" ::"r"(bar_empty) : "memory");
        }
        return;
    }
    const uint32_t tid = threadIdx.x;
    const uint32_t t = blockIdx.x * CLT256V2_B + tid;
    uint32_t *const st = (uint32_t *)smem4 + tid;
    for (uint32_t p = 0; p < CLT256V2_NSLOTS; p++) st[p * CLT256V2_B] = planes[(size_t)CLT256V2_SLOT_BIT[p] * T + t];
    const uint32_t stage_a = stage_base + tid * 4u, stage_b = stage_base + (tid ^ 16u) * 4u;
#define CLT256V2_SECTION_RESIDENT_DECL
#include "clt256_v2_gen.inc"
#undef CLT256V2_SECTION_RESIDENT_DECL

#pragma unroll 1
    for (uint32_t it = 0; it < L; it++) {
#define CLT256V2_SECTION_W_BODY
#include "clt256_v2_gen.inc"
#undef CLT256V2_SECTION_W_BODY
    }
}
#pragma pop_macro("CLT256V2_W_END")
#pragma pop_macro("CLT256V2_W_BEGIN")
#pragma pop_macro("CLT256V2_W_DRAIN")

// ---------------------------------------------------------------------------
// clt256_words7_kernel (knob words_variant = 7): the stage lives in TENSOR MEMORY.  A thread's un-sliced clock is
// 256 words; TMEM gives every lane of the four compute warps 512 private columns (128 lanes x 512 columns x 4 B =
// 256 KB per SM, reachable only with tcgen05.ld / tcgen05.st), i.e. exactly two clocks per thread -- the double
// buffer that shared memory has no room for (state 76 KB + one 128 KB stage).  Warps 0-3 run the clock and the
// eight 32x32 transposes as in words2, but a transposed word leaves its 32 registers with ONE
// tcgen05.st.32x32b.x32 (column = buffer*256 + word*32 + stream) instead of 32 st.shared; warps 4-7 sit on the same
// TMEM lane quarters (warp id mod 4), read eight streams x eight words at a time (tcgen05.ld.32x32b.x8) and store
// the 32-byte clocks.  The drain of clock `it` overlaps the WHOLE clock it+1 (words6 could only use its recurrence
// phase), and the 512 shared-memory accesses per thread and clock of the stage are gone from the LSU pipe.
// Hand-off per warp pair and buffer: mbarrier full / empty (32 arrivals each), tcgen05 fences around them.
// ---------------------------------------------------------------------------
#define CLT256V7_CTRL_BYTES 256
#define CLT256V7_SMEM (CLT256V2_STATE_BYTES + CLT256V7_CTRL_BYTES)

__device__ __forceinline__ void mbar_init(uint32_t a, uint32_t count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(a), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t a)
{
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(a) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t a, uint32_t parity)
{
    asm volatile("{\n"
                 ".reg .pred p;\n"
                 "W7_WAIT:\n"
                 "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
                 "@p bra W7_DONE;\n"
                 "bra W7_WAIT;\n"
                 "W7_DONE:\n"
                 "}" ::"r"(a), "r"(parity) : "memory");
}
__device__ __forceinline__ void tmem_st32(uint32_t taddr, const uint32_t (&r)[32])
{
    asm volatile("tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, "
                 "%15, %16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31, %32};" ::"r"(taddr),
                 "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]), "r"(r[9]),
                 "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15]), "r"(r[16]), "r"(r[17]), "r"(r[18]),
                 "r"(r[19]), "r"(r[20]), "r"(r[21]), "r"(r[22]), "r"(r[23]), "r"(r[24]), "r"(r[25]), "r"(r[26]), "r"(r[27]),
                 "r"(r[28]), "r"(r[29]), "r"(r[30]), "r"(r[31]) : "memory");
}
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t (&r)[32])
{
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, "
                 "%15, %16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
                   "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]),
                   "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]),
                   "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
                 : "r"(taddr) : "memory");
}

__device__ __forceinline__ void tmem_ld8(uint32_t taddr, uint32_t (&r)[8])
{
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7])
                 : "r"(taddr) : "memory");
}

#pragma push_macro("CLT256V2_W_DRAIN")
#pragma push_macro("CLT256V2_W_BEGIN")
#pragma push_macro("CLT256V2_W_EMIT")
#pragma push_macro("CLT256V2_W_END")
#undef CLT256V2_W_DRAIN
#undef CLT256V2_W_BEGIN
#undef CLT256V2_W_EMIT
#undef CLT256V2_W_END
// W7_STRAIGHT: the eight words of a clock as straight-line code (the cases fall through, each with its own transpose
// and tcgen05.st) instead of a loop over a switch that first copies the 32 source planes of the word into tw[]
#define CLT256V2_W_EMIT(w, r0, r1, r2, r3, r4, r5, r6, r7, r8, r9, r10, r11, r12, r13, r14, r15, r16, r17, r18, r19, \
                        r20, r21, r22, r23, r24, r25, r26, r27, r28, r29, r30, r31)                                  \
        case w: {                                                                                                    \
            const uint32_t src[32] = {r0, r1, r2, r3, r4, r5, r6, r7, r8, r9, r10, r11, r12, r13, r14, r15, r16,     \
                                      r17, r18, r19, r20, r21, r22, r23, r24, r25, r26, r27, r28, r29, r30, r31};    \
            _Pragma("unroll") for (int i = 0; i < 32; i++) tw[i] = src[i];                                           \
            if (W7_STRAIGHT) {                                                                                       \
                transpose32_fast(tw);                                                                                \
                tmem_st32(taddr + b_ * 256u + (w) * 32u, tw);                                                        \
            }                                                                                                        \
        }                                                                                                            \
        if (!W7_STRAIGHT) break;
#define CLT256V2_W_DRAIN(j)
#define CLT256V2_W_BEGIN                                                                                             \
    const uint32_t b_ = it & 1u, k_ = it >> 1;                                                                       \
    if (k_ != 0u) {                      /* the drain warp has read use k_-1 of this buffer */                        \
        mbar_wait(bar_empty + 8u * b_, (k_ - 1u) & 1u);                                                              \
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");                                              \
    }                                                                                                                \
    _Pragma("unroll 1") for (uint32_t w_ = 0; w_ < (W7_STRAIGHT ? 1u : 8u); w_++) {                                  \
        uint32_t tw[32];                                                                                             \
        switch (W7_STRAIGHT ? 0u : w_) {
#define CLT256V2_W_END                                                                                               \
        }                                                                                                            \
        if (W7_BYTES) {      /* sub-byte stages here; column 4*jl + c holds row jl + 8c for the drain's byte stages */ \
            transpose32_bit_stages<!W7_STAGE1_DRAIN>(tw);                                                            \
            uint32_t tp[32];                                                                                         \
            _Pragma("unroll") for (int cc = 0; cc < 32; cc++) tp[cc] = tw[(cc >> 2) + 8 * (cc & 3)];                 \
            tmem_st32(taddr + b_ * 256u + w_ * 32u, tp);                                                             \
        } else if (!W7_STRAIGHT) {                                                                                   \
            if (!W7_SPLIT || w_ < 4u) transpose32_fast<W7_FMA_SHR>(tw);   /* SPLIT: words 4-7 go to TMEM bit-sliced */ \
            tmem_st32(taddr + b_ * 256u + w_ * 32u, tw);                                                             \
        }                                                                                                            \
    }                                                                                                                \
    asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");                                                     \
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");                                                 \
    mbar_arrive(bar_full + 8u * b_);

// POLICY2 = 0 (knob words_variant = 8) stores every sector with evict_last, as words2 does.
// SPLIT = 1: the compute warps transpose only words 0-3 of a clock and store the planes of words 4-7 as they are; the
// drain warps transpose those four words themselves (tcgen05.ld.x32 + 32x32 register transposes), so that the two warps
// of a scheduler carry about the same number of instructions per clock (2350 / 2000 instead of 4500 / 600).
template <int POLICY2, int SPLIT>
__global__ void __launch_bounds__(2 * CLT256V2_B, 1)
clt256_words7_kernel(const uint32_t *__restrict__ planes, uint32_t T, uint32_t L, uint32_t *__restrict__ out)
{
    constexpr bool W7_SPLIT = SPLIT == 1;
    constexpr bool W7_STRAIGHT = SPLIT == 3;      // knob words_variant = 12
    constexpr bool W7_BYTES = SPLIT == 4 || SPLIT == 5;   // the byte-granular transpose stages run on the drain warps
    constexpr bool W7_STAGE1_DRAIN = SPLIT == 5;          // ... and the last sub-byte stage (rows k, k+1) as well
    constexpr bool W7_FMA_SHR = SPLIT == 2;       // knob words_variant = 11: right shifts of the transposes on the FMA pipe
    extern __shared__ uint4 smem4[];
    const uint32_t ctrl = (uint32_t)__cvta_generic_to_shared((char *)smem4 + CLT256V2_STATE_BYTES);
    const uint32_t warp = threadIdx.x >> 5, pair = warp & 3u, lane = threadIdx.x & 31u;
    if (warp == 4u) {                    // one warp allocates all 512 columns (one CTA per SM)
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(ctrl + 128u) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    if (threadIdx.x == 0)
        for (uint32_t i = 0; i < 16u; i++) mbar_init(ctrl + 8u * i, 32u);
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tbase = *(const volatile uint32_t *)((const char *)smem4 + CLT256V2_STATE_BYTES + 128);
    const uint32_t taddr = tbase + ((pair * 32u) << 16);              // this warp's lane quarter, column 0
    const uint32_t bar_full = ctrl + pair * 32u, bar_empty = bar_full + 16u;   // [buffer] 8 bytes apart
    if (warp >= 4u) {
        // lanes 2q, 2q+1 swap halves by shuffle so that a warp store covers 16 complete sectors (words2's store side; every
        // lane storing its own thread's two halves -- 32 half sectors per warp store -- ran at HALF the speed: 434 against
        // 892 GB/s on 37 CTAs): the even lane of a pair keeps the low halves (words 0-3) of both threads, the odd lane the
        // high ones
        const uint32_t odd = lane & 1u;
        uint32_t *const opa = out + (size_t)(blockIdx.x * CLT256V2_B + pair * 32u + (lane & ~1u)) * 32u * L * 8u + 4u * odd;
        uint32_t *const opb = opa + (size_t)32u * L * 8u;                 // the odd thread's streams
        // a stream's 128-byte line fills over four clocks: the first three sectors ask L2 to keep the line
        // (evict_last), the completing one releases it (evict_first), so only partial lines compete for residency
        uint64_t pol_keep, pol_done;
        asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(pol_keep));
        asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol_done));
#pragma unroll 1
        for (uint32_t it = 0; it < L; it++) {
            const uint32_t b = it & 1u, k = it >> 1;
            mbar_wait(bar_full + 8u * b, k & 1u);
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            if (W7_SPLIT) {
                uint32_t tq[4][32];                                    // words 4-7: planes in, [word][stream] out
#pragma unroll
                for (uint32_t w = 0; w < 4u; w++) tmem_ld32(taddr + b * 256u + 128u + w * 32u, tq[w]);
                asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
                for (uint32_t w = 0; w < 4u; w++) transpose32_fast(tq[w]);
#pragma unroll
                for (uint32_t jb = 0; jb < 32u; jb += 8u) {
                    uint32_t v[4][8];                                  // words 0-3 of streams jb .. jb+7
#pragma unroll
                    for (uint32_t w = 0; w < 4u; w++) tmem_ld8(taddr + b * 256u + w * 32u + jb, v[w]);
                    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
                    for (uint32_t jj = 0; jj < 8u; jj++) {
                        uint32_t keep[4], got[4];
#pragma unroll
                        for (uint32_t i = 0; i < 4u; i++) {
                            keep[i] = odd ? tq[i][jb + jj] : v[i][jj];
                            got[i] = __shfl_xor_sync(0xffffffffu, odd ? v[i][jj] : tq[i][jb + jj], 1);
                        }
                        const uint32_t sec = (jb + jj) * L + it;
                        const size_t off = (size_t)sec * 8u;
                        const uint64_t pol = POLICY2 && (sec & 3u) == 3u ? pol_done : pol_keep;
                        st_policy((uint4 *)(opa + off), odd ? make_uint4(got[0], got[1], got[2], got[3])
                                                            : make_uint4(keep[0], keep[1], keep[2], keep[3]), pol);
                        st_policy((uint4 *)(opb + off), odd ? make_uint4(keep[0], keep[1], keep[2], keep[3])
                                                            : make_uint4(got[0], got[1], got[2], got[3]), pol);
                    }
                }
            } else
#pragma unroll 1
            for (uint32_t jb = 0; jb < 32u; jb += 8u) {
                uint32_t v[8][8];
#pragma unroll
                for (uint32_t w = 0; w < 8u; w++) tmem_ld8(taddr + b * 256u + w * 32u + jb, v[w]);
                asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
                if (W7_BYTES) {
#pragma unroll
                    for (uint32_t w = 0; w < 8u; w++) {
                        if (W7_STAGE1_DRAIN) {
#pragma unroll
                            for (uint32_t cq = 0; cq < 4u; cq++) bit_stage1_pair(v[w][cq], v[w][4 + cq]);
                        }
                        byte4x4_transpose(v[w][0], v[w][1], v[w][2], v[w][3]);
                        byte4x4_transpose(v[w][4], v[w][5], v[w][6], v[w][7]);
                    }
                }
#pragma unroll
                for (uint32_t jj = 0; jj < 8u; jj++) {
                        uint32_t keep[4], got[4];
#pragma unroll
                        for (uint32_t i = 0; i < 4u; i++) {
                            keep[i] = odd ? v[4 + i][jj] : v[i][jj];
                            got[i] = __shfl_xor_sync(0xffffffffu, odd ? v[i][jj] : v[4 + i][jj], 1);
                        }
                        // BYTES: this batch holds streams 2m + {0, 1} + 8c, m = jb / 8
                        const uint32_t sj = W7_BYTES ? (jb >> 2) + (jj >> 2) + 8u * (jj & 3u) : jb + jj;
                        const uint32_t sec = sj * L + it;                 // sector number within the thread's 32 streams
                        const size_t off = (size_t)sec * 8u;
                        // (releasing HALF lines -- evict_first on every odd sector -- measured 2497 against 2834 GB/s)
                        const uint64_t pol = POLICY2 && (sec & 3u) == 3u ? pol_done : pol_keep;
                        // first store: the even thread's stream (even lane: its own low half, odd lane: the received high half)
                        st_policy((uint4 *)(opa + off), odd ? make_uint4(got[0], got[1], got[2], got[3])
                                                            : make_uint4(keep[0], keep[1], keep[2], keep[3]), pol);
                        st_policy((uint4 *)(opb + off), odd ? make_uint4(keep[0], keep[1], keep[2], keep[3])
                                                            : make_uint4(got[0], got[1], got[2], got[3]), pol);
                }
            }
            asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
            mbar_arrive(bar_empty + 8u * b);
        }
    } else {
        const uint32_t tid = threadIdx.x;
        const uint32_t t = blockIdx.x * CLT256V2_B + tid;
        uint32_t *const st = (uint32_t *)smem4 + tid;
        for (uint32_t p = 0; p < CLT256V2_NSLOTS; p++) st[p * CLT256V2_B] = planes[(size_t)CLT256V2_SLOT_BIT[p] * T + t];
#define CLT256V2_SECTION_RESIDENT_DECL
#include "clt256_v2_gen.inc"
#undef CLT256V2_SECTION_RESIDENT_DECL

#pragma unroll 1
        for (uint32_t it = 0; it < L; it++) {
#define CLT256V2_SECTION_W_BODY
#include "clt256_v2_gen.inc"
#undef CLT256V2_SECTION_W_BODY
        }
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 4u) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tbase) : "memory");
}
// ---------------------------------------------------------------------------
// clt256_words10_kernel (knob words_variant = 16): words7 with TWO drain warps per compute warp and the register file
// re-divided by setmaxnreg (launch: 384 threads x 168; then the compute warpgroup 248, the two drain warpgroups 128
// each: 4 x 32 x 248 + 8 x 32 x 128 = 64 512).  Each drain warp takes two of the four 8-stream batches of a clock, so
// the drain's dependent chains (TMEM load -> byte / bit stages -> shuffle -> store) overlap twice as well.
// ---------------------------------------------------------------------------
__global__ void __launch_bounds__(3 * CLT256V2_B, 1)
clt256_words10_kernel(const uint32_t *__restrict__ planes, uint32_t T, uint32_t L, uint32_t *__restrict__ out)
{
    constexpr bool W7_SPLIT = false, W7_STRAIGHT = false, W7_BYTES = true, W7_STAGE1_DRAIN = true, W7_FMA_SHR = false;
    (void)W7_SPLIT; (void)W7_FMA_SHR;
    extern __shared__ uint4 smem4[];
    const uint32_t ctrl = (uint32_t)__cvta_generic_to_shared((char *)smem4 + CLT256V2_STATE_BYTES);
    const uint32_t warp = threadIdx.x >> 5, pair = warp & 3u, lane = threadIdx.x & 31u;
    if (warp == 4u) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(ctrl + 128u) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    if (threadIdx.x == 0)
        for (uint32_t i = 0; i < 16u; i++) mbar_init(ctrl + 8u * i, (i & 2u) ? 64u : 32u);   // full: 32 arrivals, empty: 64
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tbase = *(const volatile uint32_t *)((const char *)smem4 + CLT256V2_STATE_BYTES + 128);
    const uint32_t taddr = tbase + ((pair * 32u) << 16);
    const uint32_t bar_full = ctrl + pair * 32u, bar_empty = bar_full + 16u;
    if (warp >= 4u) {
        asm volatile("setmaxnreg.dec.sync.aligned.u32 128;");
        const uint32_t half = (warp >> 2) - 1u;                          // 0: batches 0, 1; 1: batches 2, 3
        const uint32_t odd = lane & 1u;
        uint32_t *const opa = out + (size_t)(blockIdx.x * CLT256V2_B + pair * 32u + (lane & ~1u)) * 32u * L * 8u + 4u * odd;
        uint32_t *const opb = opa + (size_t)32u * L * 8u;
        uint64_t pol_keep, pol_done;
        asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(pol_keep));
        asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol_done));
#pragma unroll 1
        for (uint32_t it = 0; it < L; it++) {
            const uint32_t b = it & 1u, k = it >> 1;
            mbar_wait(bar_full + 8u * b, k & 1u);
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
#pragma unroll 1
            for (uint32_t jb = half * 16u; jb < half * 16u + 16u; jb += 8u) {
                uint32_t v[8][8];
#pragma unroll
                for (uint32_t w = 0; w < 8u; w++) tmem_ld8(taddr + b * 256u + w * 32u + jb, v[w]);
                asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
                for (uint32_t w = 0; w < 8u; w++) {
#pragma unroll
                    for (uint32_t cq = 0; cq < 4u; cq++) bit_stage1_pair(v[w][cq], v[w][4 + cq]);
                    byte4x4_transpose(v[w][0], v[w][1], v[w][2], v[w][3]);
                    byte4x4_transpose(v[w][4], v[w][5], v[w][6], v[w][7]);
                }
#pragma unroll
                for (uint32_t jj = 0; jj < 8u; jj++) {
                    uint32_t keep[4], got[4];
#pragma unroll
                    for (uint32_t i = 0; i < 4u; i++) {
                        keep[i] = odd ? v[4 + i][jj] : v[i][jj];
                        got[i] = __shfl_xor_sync(0xffffffffu, odd ? v[i][jj] : v[4 + i][jj], 1);
                    }
                    const uint32_t sj = (jb >> 2) + (jj >> 2) + 8u * (jj & 3u);
                    const uint32_t sec = sj * L + it;
                    const size_t off = (size_t)sec * 8u;
                    const uint64_t pol = (sec & 3u) == 3u ? pol_done : pol_keep;
                    st_policy((uint4 *)(opa + off), odd ? make_uint4(got[0], got[1], got[2], got[3])
                                                        : make_uint4(keep[0], keep[1], keep[2], keep[3]), pol);
                    st_policy((uint4 *)(opb + off), odd ? make_uint4(keep[0], keep[1], keep[2], keep[3])
                                                        : make_uint4(got[0], got[1], got[2], got[3]), pol);
                }
            }
            asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
            mbar_arrive(bar_empty + 8u * b);
        }
    } else {
        asm volatile("setmaxnreg.inc.sync.aligned.u32 248;");
        const uint32_t tid = threadIdx.x;
        const uint32_t t = blockIdx.x * CLT256V2_B + tid;
        uint32_t *const st = (uint32_t *)smem4 + tid;
        for (uint32_t p = 0; p < CLT256V2_NSLOTS; p++) st[p * CLT256V2_B] = planes[(size_t)CLT256V2_SLOT_BIT[p] * T + t];
#define CLT256V2_SECTION_RESIDENT_DECL
#include "clt256_v2_gen.inc"
#undef CLT256V2_SECTION_RESIDENT_DECL

#pragma unroll 1
        for (uint32_t it = 0; it < L; it++) {
#define CLT256V2_SECTION_W_BODY
#include "clt256_v2_gen.inc"
#undef CLT256V2_SECTION_W_BODY
        }
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 4u) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tbase) : "memory");
}
#pragma pop_macro("CLT256V2_W_END")
#pragma pop_macro("CLT256V2_W_EMIT")
#pragma pop_macro("CLT256V2_W_BEGIN")
#pragma pop_macro("CLT256V2_W_DRAIN")

// ---------------------------------------------------------------------------
// clt256_words8_kernel (knob words_variant = 9): words7 with the UN-SLICING moved to the drain warps as well.  In
// words7 a scheduler holds one compute warp with 4500 instructions per clock (recurrence 1600 + transposes 2400 +
// the rest) and one drain warp with 600: the compute warp alone cannot issue more than every other cycle (53 % of the
// slots).  Here the compute warps hand the NEW STATE to tensor memory as it is, bit-sliced (the 32 planes of word w
// with one tcgen05.st.x32, column = buffer*256 + word*32 + plane), and a drain warp reads them back four times, once
// per group of 8 streams: byte g of the 32 planes of a word gathered with 24 PRMT, an 8x8 bit transpose in three
// shift/LOP3 stages (unslice8 of words4), eight words -> 8 streams x 32 bytes in 64 registers, lane pairs, stores.
// Per clock about 1900 instructions on the compute warp and 3300 on the drain warp of the same scheduler.
// ---------------------------------------------------------------------------
#pragma push_macro("CLT256V2_W_DRAIN")
#pragma push_macro("CLT256V2_W_BEGIN")
#pragma push_macro("CLT256V2_W_END")
#undef CLT256V2_W_DRAIN
#undef CLT256V2_W_BEGIN
#undef CLT256V2_W_END
#define CLT256V2_W_DRAIN(j)
#define CLT256V2_W_BEGIN                                                                                             \
    const uint32_t b_ = it & 1u, k_ = it >> 1;                                                                       \
    if (k_ != 0u) {                      /* the drain warp has read use k_-1 of this buffer */                        \
        mbar_wait(bar_empty + 8u * b_, (k_ - 1u) & 1u);                                                              \
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");                                              \
    }                                                                                                                \
    _Pragma("unroll 1") for (uint32_t w_ = 0; w_ < 8u; w_++) {                                                       \
        uint32_t tw[32];                                                                                             \
        switch (w_) {
#define CLT256V2_W_END                                                                                               \
        }                                                                                                            \
        tmem_st32(taddr + b_ * 256u + w_ * 32u, tw);                                                                 \
    }                                                                                                                \
    asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");                                                     \
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");                                                 \
    mbar_arrive(bar_full + 8u * b_);

__global__ void __launch_bounds__(2 * CLT256V2_B, 1)
clt256_words8_kernel(const uint32_t *__restrict__ planes, uint32_t T, uint32_t L, uint32_t *__restrict__ out)
{
    extern __shared__ uint4 smem4[];
    const uint32_t ctrl = (uint32_t)__cvta_generic_to_shared((char *)smem4 + CLT256V2_STATE_BYTES);
    const uint32_t warp = threadIdx.x >> 5, pair = warp & 3u, lane = threadIdx.x & 31u;
    if (warp == 4u) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(ctrl + 128u) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    if (threadIdx.x == 0)
        for (uint32_t i = 0; i < 16u; i++) mbar_init(ctrl + 8u * i, 32u);
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tbase = *(const volatile uint32_t *)((const char *)smem4 + CLT256V2_STATE_BYTES + 128);
    const uint32_t taddr = tbase + ((pair * 32u) << 16);
    const uint32_t bar_full = ctrl + pair * 32u, bar_empty = bar_full + 16u;
    if (warp >= 4u) {
        const uint32_t odd = lane & 1u;
        uint32_t *const opa = out + (size_t)(blockIdx.x * CLT256V2_B + pair * 32u + (lane & ~1u)) * 32u * L * 8u + 4u * odd;
        uint32_t *const opb = opa + (size_t)32u * L * 8u;
        uint64_t pol_keep, pol_done;
        asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(pol_keep));
        asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol_done));
#pragma unroll 1
        for (uint32_t it = 0; it < L; it++) {
            const uint32_t b = it & 1u, k = it >> 1;
            mbar_wait(bar_full + 8u * b, k & 1u);
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            // software pipeline over the 16 (group, word pair) steps of a clock: the planes of the next step are on their
            // way from TMEM while this step's are un-sliced (tcgen05.wait::ld waits for every load issued, so the next
            // load is issued right after the wait)
            uint32_t ta[2][32], tb[2][32];
            const uint32_t tcol = taddr + b * 256u;
            tmem_ld32(tcol, ta[0]);
            tmem_ld32(tcol + 32u, tb[0]);
#pragma unroll 1
            for (uint32_t g = 0; g < 4u; g++) {
                const uint32_t sel2 = g | ((4u + g) << 4);
                uint32_t v[8][8];                                      // [word][stream of the group]
#pragma unroll
                for (uint32_t p = 0; p < 4u; p++) {                    // words 2p, 2p+1
                    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
                    if (p < 3u || g < 3u) {
                        const uint32_t nw = p < 3u ? 2u * (p + 1u) : 0u;
                        tmem_ld32(tcol + nw * 32u, ta[(p + 1u) & 1u]);
                        tmem_ld32(tcol + nw * 32u + 32u, tb[(p + 1u) & 1u]);
                    }
                    unslice8(ta[p & 1u], sel2, v[2u * p]);
                    unslice8(tb[p & 1u], sel2, v[2u * p + 1u]);
                }
#pragma unroll
                for (uint32_t jj = 0; jj < 8u; jj++) {
                    uint32_t keep[4], got[4];
#pragma unroll
                    for (uint32_t i = 0; i < 4u; i++) {
                        keep[i] = odd ? v[4 + i][jj] : v[i][jj];
                        got[i] = __shfl_xor_sync(0xffffffffu, odd ? v[i][jj] : v[4 + i][jj], 1);
                    }
                    const uint32_t sec = (g * 8u + jj) * L + it;
                    const size_t off = (size_t)sec * 8u;
                    const uint64_t pol = (sec & 3u) == 3u ? pol_done : pol_keep;
                    st_policy((uint4 *)(opa + off), odd ? make_uint4(got[0], got[1], got[2], got[3])
                                                        : make_uint4(keep[0], keep[1], keep[2], keep[3]), pol);
                    st_policy((uint4 *)(opb + off), odd ? make_uint4(keep[0], keep[1], keep[2], keep[3])
                                                        : make_uint4(got[0], got[1], got[2], got[3]), pol);
                }
            }
            asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
            mbar_arrive(bar_empty + 8u * b);
        }
    } else {
        const uint32_t tid = threadIdx.x;
        const uint32_t t = blockIdx.x * CLT256V2_B + tid;
        uint32_t *const st = (uint32_t *)smem4 + tid;
        for (uint32_t p = 0; p < CLT256V2_NSLOTS; p++) st[p * CLT256V2_B] = planes[(size_t)CLT256V2_SLOT_BIT[p] * T + t];
#define CLT256V2_SECTION_RESIDENT_DECL
#include "clt256_v2_gen.inc"
#undef CLT256V2_SECTION_RESIDENT_DECL

#pragma unroll 1
        for (uint32_t it = 0; it < L; it++) {
#define CLT256V2_SECTION_W_BODY
#include "clt256_v2_gen.inc"
#undef CLT256V2_SECTION_W_BODY
        }
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 4u) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tbase) : "memory");
}
#pragma pop_macro("CLT256V2_W_END")
#pragma pop_macro("CLT256V2_W_BEGIN")
#pragma pop_macro("CLT256V2_W_DRAIN")

// ---------------------------------------------------------------------------
// clt256_words9_kernel (knob words_variant = 13): words7 with THREE clocks of staging -- the two tensor-memory
// buffers and the 128 KB shared-memory stage of words2 -- so that the drain warps take the clocks in PAIRS: the two
// sectors a stream gets from clocks 2k, 2k+1 are adjacent (64 bytes of one 128-byte line) and leave back to back.
// tools/probe_store_pattern.cu: 606 k streams take 2575 GB/s in 32-byte visits and 4041 GB/s in 64-byte visits.
// Clock `it` goes to buffer it % 3 (0, 1: TMEM columns 0 / 256; 2: shared memory, word [stream][word][thread]); pair k
// = buffers (2k % 3, (2k + 1) % 3) is drained while clock 2k + 2 is computed, clock 2k + 3 re-uses the first of them.
// ---------------------------------------------------------------------------
#define CLT256V9_STAGE_BYTES (32 * 8 * CLT256V2_B * 4)
#define CLT256V9_SMEM (CLT256V2_STATE_BYTES + CLT256V9_STAGE_BYTES + CLT256V7_CTRL_BYTES)

#pragma push_macro("CLT256V2_W_DRAIN")
#pragma push_macro("CLT256V2_W_BEGIN")
#pragma push_macro("CLT256V2_W_END")
#undef CLT256V2_W_DRAIN
#undef CLT256V2_W_BEGIN
#undef CLT256V2_W_END
#define CLT256V2_W_DRAIN(j)
#define CLT256V2_W_BEGIN                                                                                             \
    const uint32_t b_ = it % 3u, u_ = it / 3u;                                                                       \
    if (u_ != 0u) {                      /* the drain warp has read use u_-1 of this buffer */                        \
        mbar_wait(bar_empty + 8u * b_, (u_ - 1u) & 1u);                                                              \
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");                                              \
    }                                                                                                                \
    _Pragma("unroll 1") for (uint32_t w_ = 0; w_ < 8u; w_++) {                                                       \
        uint32_t tw[32];                                                                                             \
        switch (w_) {
#define CLT256V2_W_END                                                                                               \
        }                                                                                                            \
        transpose32_fast(tw);                                                                                        \
        if (b_ < 2u) {                                                                                               \
            tmem_st32(taddr + b_ * 256u + w_ * 32u, tw);                                                             \
        } else {                                                                                                     \
            const uint32_t sa = stage_t + w_ * (uint32_t)(CLT256V2_B * 4);                                           \
            _Pragma("unroll") for (int j = 0; j < 32; j++)                                                           \
                asm volatile("st.shared.u32 [%0], %1;" ::"r"(sa + (uint32_t)(j * 8 * CLT256V2_B * 4)), "r"(tw[j]) : "memory"); \
        }                                                                                                            \
    }                                                                                                                \
    asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");                                                     \
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");                                                 \
    mbar_arrive(bar_full + 8u * b_);                     /* release: the shared-memory stores are visible too */

// eight streams x eight words of one staged clock: from tensor memory (buf 0, 1) or the shared-memory stage (buf 2)
__device__ __forceinline__ void w9_load(uint32_t buf, uint32_t taddr, uint32_t stage_t, uint32_t jb, uint32_t (&v)[8][8])
{
    if (buf < 2u) {
#pragma unroll
        for (uint32_t w = 0; w < 8u; w++) tmem_ld8(taddr + buf * 256u + w * 32u + jb, v[w]);
    } else {
#pragma unroll
        for (uint32_t w = 0; w < 8u; w++)
#pragma unroll
            for (uint32_t jj = 0; jj < 8u; jj++)
                asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v[w][jj])
                             : "r"(stage_t + (uint32_t)((((jb + jj) * 8u + w) * CLT256V2_B) * 4u)) : "memory");
    }
}

// the lane pair's two 16-byte stores of one stream's 32-byte clock (see clt256_words7_kernel)
__device__ __forceinline__ void w9_store(const uint32_t (&v)[8][8], uint32_t jj, uint32_t odd, uint32_t *pa, uint32_t *pb,
                                         uint64_t pol)
{
    uint32_t keep[4], got[4];
#pragma unroll
    for (uint32_t i = 0; i < 4u; i++) {
        keep[i] = odd ? v[4 + i][jj] : v[i][jj];
        got[i] = __shfl_xor_sync(0xffffffffu, odd ? v[i][jj] : v[4 + i][jj], 1);
    }
    st_policy((uint4 *)pa, odd ? make_uint4(got[0], got[1], got[2], got[3]) : make_uint4(keep[0], keep[1], keep[2], keep[3]), pol);
    st_policy((uint4 *)pb, odd ? make_uint4(keep[0], keep[1], keep[2], keep[3]) : make_uint4(got[0], got[1], got[2], got[3]), pol);
}

__global__ void __launch_bounds__(2 * CLT256V2_B, 1)
clt256_words9_kernel(const uint32_t *__restrict__ planes, uint32_t T, uint32_t L, uint32_t *__restrict__ out)
{
    extern __shared__ uint4 smem4[];
    const uint32_t stage_base = (uint32_t)__cvta_generic_to_shared((char *)smem4 + CLT256V2_STATE_BYTES);
    const uint32_t ctrl = stage_base + CLT256V9_STAGE_BYTES;
    const uint32_t warp = threadIdx.x >> 5, pair = warp & 3u, lane = threadIdx.x & 31u;
    if (warp == 4u) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(ctrl + 192u) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    if (threadIdx.x == 0)
        for (uint32_t i = 0; i < 24u; i++) mbar_init(ctrl + 8u * i, 32u);     // [warp pair][full 0-2, empty 0-2]
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tbase = *(const volatile uint32_t *)((const char *)smem4 + CLT256V2_STATE_BYTES + CLT256V9_STAGE_BYTES + 192);
    const uint32_t taddr = tbase + ((pair * 32u) << 16);
    const uint32_t bar_full = ctrl + pair * 48u, bar_empty = bar_full + 24u;
    const uint32_t stage_t = stage_base + (pair * 32u + lane) * 4u;            // this thread's column of the stage
    if (warp >= 4u) {
        const uint32_t odd = lane & 1u;
        uint32_t *const opa = out + (size_t)(blockIdx.x * CLT256V2_B + pair * 32u + (lane & ~1u)) * 32u * L * 8u + 4u * odd;
        uint32_t *const opb = opa + (size_t)32u * L * 8u;
        uint64_t pol_keep, pol_done;
        asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(pol_keep));
        asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol_done));
#pragma unroll 1
        for (uint32_t it = 0; it < L; it += 2u) {
            const bool two = it + 1u < L;
            const uint32_t ba = it % 3u, bb = (it + 1u) % 3u;
            mbar_wait(bar_full + 8u * ba, (it / 3u) & 1u);
            if (two) mbar_wait(bar_full + 8u * bb, ((it + 1u) / 3u) & 1u);
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
#pragma unroll 1
            for (uint32_t jb = 0; jb < 32u; jb += 8u) {
                uint32_t va[8][8], vb[8][8];
                w9_load(ba, taddr, stage_t, jb, va);
                if (two) w9_load(bb, taddr, stage_t, jb, vb);
                asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
                for (uint32_t jj = 0; jj < 8u; jj++) {
                    const uint32_t sec = (jb + jj) * L + it;
                    const size_t off = (size_t)sec * 8u;
                    w9_store(va, jj, odd, opa + off, opb + off, (sec & 3u) == 3u ? pol_done : pol_keep);
                    if (two) w9_store(vb, jj, odd, opa + off + 8, opb + off + 8, ((sec + 1u) & 3u) == 3u ? pol_done : pol_keep);
                }
            }
            asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
            mbar_arrive(bar_empty + 8u * ba);
            if (two) mbar_arrive(bar_empty + 8u * bb);
        }
    } else {
        const uint32_t tid = threadIdx.x;
        const uint32_t t = blockIdx.x * CLT256V2_B + tid;
        uint32_t *const st = (uint32_t *)smem4 + tid;
        for (uint32_t p = 0; p < CLT256V2_NSLOTS; p++) st[p * CLT256V2_B] = planes[(size_t)CLT256V2_SLOT_BIT[p] * T + t];
#define CLT256V2_SECTION_RESIDENT_DECL
#include "clt256_v2_gen.inc"
#undef CLT256V2_SECTION_RESIDENT_DECL

#pragma unroll 1
        for (uint32_t it = 0; it < L; it++) {
#define CLT256V2_SECTION_W_BODY
#include "clt256_v2_gen.inc"
#undef CLT256V2_SECTION_W_BODY
        }
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 4u) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tbase) : "memory");
}
#pragma pop_macro("CLT256V2_W_END")
#pragma pop_macro("CLT256V2_W_BEGIN")
#pragma pop_macro("CLT256V2_W_DRAIN")

// ---------------------------------------------------------------------------
// clt256_words3_kernel (EXPERIMENT, knob "words_b192"; measured and rejected): the same clock and transposes with
// SIX warps per SM instead of four.  ncu on words2 shows neither pipe busy (ALU 34 %, LSU 36 %), so more warps
// looked like the fix; but 192 threads need 114 KB of state, which leaves room for only HALF a clock of staging
// (4 of the 8 words of every stream: 96 KB), so a clock leaves in two passes of one 16-byte piece per stream.
// Result at 2^28 steps: 494 GB/s against 1658 GB/s for words2 -- half-sector stores cost far more than the extra
// warps give, even with the pieces meeting in L2 (evict_last).  Complete 32-byte sectors per store are what matters.
// The generated body is re-instantiated with B = 192.
// ---------------------------------------------------------------------------
#pragma push_macro("CLT256V2_B")
#pragma push_macro("CLT256V2_W_EMIT")
#undef CLT256V2_B
#undef CLT256V2_W_EMIT
#pragma push_macro("CLT256V2_W_DRAIN")
#pragma push_macro("CLT256V2_W_BEGIN")
#pragma push_macro("CLT256V2_W_END")
#undef CLT256V2_W_DRAIN
#undef CLT256V2_W_BEGIN
#undef CLT256V2_W_END
#define CLT256V2_W_DRAIN(j)
#define CLT256V2_W_BEGIN
#define CLT256V2_W_END
#define CLT256V2_B 192
#define CLT256V3W_STAGE_WORDS (32 * 4 * 192)
#define CLT256V3W_STATE_BYTES (CLT256V2_NSLOTS * 192 * 4)
#define CLT256V3W_SMEM (CLT256V3W_STATE_BYTES + CLT256V3W_STAGE_WORDS * 4)

#define CLT256V2_W_EMIT(w, r0, r1, r2, r3, r4, r5, r6, r7, r8, r9, r10, r11, r12, r13, r14, r15, r16, r17, r18, r19, \
                        r20, r21, r22, r23, r24, r25, r26, r27, r28, r29, r30, r31)                                  \
    {                                                                                                                \
        uint32_t tw[32] = {r0, r1, r2, r3, r4, r5, r6, r7, r8, r9, r10, r11, r12, r13, r14, r15, r16, r17, r18,       \
                           r19, r20, r21, r22, r23, r24, r25, r26, r27, r28, r29, r30, r31};                          \
        transpose32_fast(tw);                                                                                        \
        _Pragma("unroll") for (int j = 0; j < 32; j++)                                                               \
            asm volatile("st.shared.u32 [%0], %1;" ::"r"(stage_a + (uint32_t)((j * 4 + ((w) & 3)) * 192 * 4)), "r"(tw[j])); \
        if (((w) & 3) == 3) {                                                                                        \
            uint32_t *const o32 = outp + (size_t)it * 8u + ((w) >> 2) * 4u;                                          \
            _Pragma("unroll 8") for (int j = 0; j < 32; j++) {                                                       \
                uint32_t v[4];                                                                                       \
                _Pragma("unroll") for (int i = 0; i < 4; i++)                                                        \
                    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v[i]) : "r"(stage_a + (uint32_t)((j * 4 + i) * 192 * 4))); \
                CLT256V2W_STORE((uint4 *)(o32 + (size_t)j * L * 8u), make_uint4(v[0], v[1], v[2], v[3]));             \
            }                                                                                                        \
        }                                                                                                            \
    }

__global__ void __launch_bounds__(192, 1)
clt256_words3_kernel(const uint32_t *__restrict__ planes, uint32_t T, uint32_t L, uint32_t *__restrict__ out)
{
    extern __shared__ uint4 smem4[];
    const uint32_t tid = threadIdx.x;
    const uint32_t t = blockIdx.x * 192u + tid;
    uint32_t *const st = (uint32_t *)smem4 + tid;
    for (uint32_t p = 0; p < CLT256V2_NSLOTS; p++) st[p * 192u] = planes[(size_t)CLT256V2_SLOT_BIT[p] * T + t];
    const uint32_t stage_a = (uint32_t)__cvta_generic_to_shared((char *)smem4 + CLT256V3W_STATE_BYTES) + tid * 4u;
    uint32_t *const outp = out + (size_t)t * 32u * L * 8u;   // stream j, clock it: outp + (j*L + it)*8
#define CLT256V2_SECTION_RESIDENT_DECL
#include "clt256_v2_gen.inc"
#undef CLT256V2_SECTION_RESIDENT_DECL

#pragma unroll 1
    for (uint32_t it = 0; it < L; it++) {
#define CLT256V2_SECTION_W_BODY
#include "clt256_v2_gen.inc"
#undef CLT256V2_SECTION_W_BODY
    }
}
#pragma pop_macro("CLT256V2_W_END")
#pragma pop_macro("CLT256V2_W_BEGIN")
#pragma pop_macro("CLT256V2_W_DRAIN")
#pragma pop_macro("CLT256V2_W_EMIT")
#pragma pop_macro("CLT256V2_B")

// kind: 1 = int8 samples, 2 = raw words, 3 = int16 scaled samples
int bbb_clt_spec_stream_ok(const bbb_ctx *c, int kind)
{
    (void)kind;
    return c->shipped == 256 && !c->tune_no_spec;
}

int bbb_clt_spec_stream_launch(bbb_ctx *c, int kind, uint32_t blocks, uint32_t L, const uint32_t *planes, void *out,
                               uint32_t noise_var)
{
    if ((kind == 1 || kind == 3) && c->tune_spec_variant != 1) {
        static_assert(CLT256V2_CTAS_PER_SM * (CLT256V2_STATE_BYTES + 1024) <= 228 * 1024, "shared memory budget");
        const uint32_t T = blocks * CLT256V2_B;
        uint32_t *slab = (uint32_t *)c->scratch(11, (size_t)CLT256V2_S8_K * 8 * T * 4);
        if (!slab) return BBB_ECUDA;
        BBB_CUDA(c, cudaFuncSetAttribute(clt256_s8v2_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         CLT256V2_STATE_BYTES));
        clt256_s8v2_kernel<<<blocks, CLT256V2_B, CLT256V2_STATE_BYTES, c->stream>>>(
            planes, T, L, (int8_t *)out, kind == 1 ? -1 : (int32_t)noise_var, slab);
        c->last_kernel = kind == 1 ? "clt256_s8v2" : "clt256_s16v2";
    } else if (kind == 1) {
        static_assert(CLT256S8_SMEM <= 227 * 1024, "shared memory budget");
        BBB_CUDA(c, cudaFuncSetAttribute(clt256_s8_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, CLT256S8_SMEM));
        clt256_s8_kernel<<<blocks, CLT256_B, CLT256S8_SMEM, c->stream>>>(planes, blocks * CLT256_B, L, (int8_t *)out, -1);
        c->last_kernel = "clt256_s8";
    } else if (kind == 3) {
        BBB_CUDA(c, cudaFuncSetAttribute(clt256_s8_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, CLT256S8_SMEM));
        clt256_s8_kernel<<<blocks, CLT256_B, CLT256S8_SMEM, c->stream>>>(planes, blocks * CLT256_B, L, (int8_t *)out,
                                                                             (int32_t)noise_var);
        c->last_kernel = "clt256_s16";
    } else if (c->tune_spec_variant != 1 && c->tune_words_variant == 16) {
        BBB_CUDA(c, cudaFuncSetAttribute(clt256_words10_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, CLT256V7_SMEM));
        clt256_words10_kernel<<<blocks, 3 * CLT256V2_B, CLT256V7_SMEM, c->stream>>>(planes, blocks * CLT256V2_B, L, (uint32_t *)out);
        c->last_kernel = "clt256_words10";
    } else if (c->tune_spec_variant != 1 && c->tune_words_variant == 13) {
        static_assert(CLT256V9_SMEM <= 227 * 1024, "shared memory budget");
        BBB_CUDA(c, cudaFuncSetAttribute(clt256_words9_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, CLT256V9_SMEM));
        clt256_words9_kernel<<<blocks, 2 * CLT256V2_B, CLT256V9_SMEM, c->stream>>>(planes, blocks * CLT256V2_B, L, (uint32_t *)out);
        c->last_kernel = "clt256_words9";
    } else if (c->tune_spec_variant != 1 && c->tune_words_variant == 9) {
        BBB_CUDA(c, cudaFuncSetAttribute(clt256_words8_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, CLT256V7_SMEM));
        clt256_words8_kernel<<<blocks, 2 * CLT256V2_B, CLT256V7_SMEM, c->stream>>>(planes, blocks * CLT256V2_B, L, (uint32_t *)out);
        c->last_kernel = "clt256_words8";
    } else if (c->tune_spec_variant != 1 && c->tune_words_variant == 6) {
        BBB_CUDA(c, cudaFuncSetAttribute(clt256_words6_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, CLT256V2W_SMEM));
        clt256_words6_kernel<<<blocks, 2 * CLT256V2_B, CLT256V2W_SMEM, c->stream>>>(planes, blocks * CLT256V2_B, L, (uint32_t *)out);
        c->last_kernel = "clt256_words6";
    } else if (c->tune_spec_variant != 1 && c->tune_words_variant == 5) {
        const uint32_t T = blocks * CLT256V2_B;
        uint32_t *slab = (uint32_t *)c->scratch(11, (size_t)256 * T * 4);
        if (!slab) return BBB_ECUDA;
        BBB_CUDA(c, cudaFuncSetAttribute(clt256_words5_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, CLT256V2_STATE_BYTES));
        clt256_words5_kernel<<<blocks, CLT256V2_B, CLT256V2_STATE_BYTES, c->stream>>>(planes, T, L, (uint32_t *)out, slab);
        c->last_kernel = "clt256_words5";
    } else if (c->tune_spec_variant != 1 && c->tune_words_variant == 4) {
        static_assert(2 * (CLT256V4W_SMEM + 1024) <= 228 * 1024, "shared memory budget");
        BBB_CUDA(c, cudaFuncSetAttribute(clt256_words4_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, CLT256V4W_SMEM));
        clt256_words4_kernel<<<blocks, CLT256V2_B, CLT256V4W_SMEM, c->stream>>>(planes, blocks * CLT256V2_B, L, (uint32_t *)out);
        c->last_kernel = "clt256_words4";
    } else if (c->tune_spec_variant != 1 && c->tune_words_b192) {
        static_assert(CLT256V3W_SMEM <= 227 * 1024, "shared memory budget");
        BBB_CUDA(c, cudaFuncSetAttribute(clt256_words3_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, CLT256V3W_SMEM));
        clt256_words3_kernel<<<blocks, 192, CLT256V3W_SMEM, c->stream>>>(planes, blocks * 192, L, (uint32_t *)out);
        c->last_kernel = "clt256_words3";
    } else if (c->tune_spec_variant != 1 && c->tune_words_variant != 2) {
        auto kern = c->tune_words_variant == 8 ? clt256_words7_kernel<0, 0>
                    : c->tune_words_variant == 10 ? clt256_words7_kernel<1, 1>
                    : c->tune_words_variant == 11 ? clt256_words7_kernel<1, 2>
                    : c->tune_words_variant == 12 ? clt256_words7_kernel<1, 3>
                    : c->tune_words_variant == 7 ? clt256_words7_kernel<1, 0>
                    : c->tune_words_variant == 14 ? clt256_words7_kernel<1, 4> : clt256_words7_kernel<1, 5>;   // 0 and 15
        BBB_CUDA(c, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, CLT256V7_SMEM));
        kern<<<blocks, 2 * CLT256V2_B, CLT256V7_SMEM, c->stream>>>(planes, blocks * CLT256V2_B, L, (uint32_t *)out);
        c->last_kernel = "clt256_words7";
    } else if (c->tune_spec_variant != 1) {
        static_assert(CLT256V2W_SMEM <= 227 * 1024, "shared memory budget");
        BBB_CUDA(c, cudaFuncSetAttribute(clt256_words2_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, CLT256V2W_SMEM));
        clt256_words2_kernel<<<blocks, CLT256V2_B, CLT256V2W_SMEM, c->stream>>>(planes, blocks * CLT256V2_B, L, (uint32_t *)out);
        c->last_kernel = "clt256_words2";
    } else {
        BBB_CUDA(c, cudaFuncSetAttribute(clt256_words_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         CLT256_STATE_BYTES));
        clt256_words_kernel<<<blocks, CLT256_B, CLT256_STATE_BYTES, c->stream>>>(planes, blocks * CLT256_B, L,
                                                                                 (uint32_t *)out);
        c->last_kernel = "clt256_words";
    }
    c->last_launches++;
    BBB_CUDA(c, cudaGetLastError());
    return BBB_OK;
}
