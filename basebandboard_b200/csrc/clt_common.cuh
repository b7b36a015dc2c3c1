// clt_common.cuh -- device helpers shared by the CLT / LUTOPT kernels.
#pragma once
#include <cstdint>

struct SeedStart {
    uint32_t w[16];   // level start state, 32-bit little-endian words
};

// One level of the stream decomposition handed to a specialised kernel.
struct SpecLevel {
    uint32_t T;            // main-kernel threads (32 streams each)
    uint32_t logL;         // every stream runs 2^logL clocks
    uint64_t nvalid;       // streams s < nvalid are real
    uint64_t out_base;     // sample offset of stream 0 in the output arrays
    uint32_t model, noise_var;
    const uint32_t *planes;        // [n][T] bit-sliced start states
    unsigned long long *hist;      // n+1 bins or null
    int8_t *out8;
    int16_t *out16;
    uint32_t *out_words;
};

// carry-save adder: (carry, sum) = a + b + c, bitwise over 32 streams (2 LOP3)
__device__ __forceinline__ void csa(uint32_t &carry, uint32_t &sum, uint32_t a, uint32_t b, uint32_t c)
{
    const uint32_t s = a ^ b ^ c;
    carry = (a & b) | (a & c) | (b & c);
    sum = s;
}

// In-place transpose of a 32x32 bit matrix: on return bit k of a[j] is what bit j
// of a[k] was (LSB = column 0).  5 butterfly stages.
__device__ __forceinline__ void transpose32(uint32_t (&a)[32])
{
#pragma unroll
    for (int s = 0; s < 5; s++) {
        const int j = 16 >> s;
        const uint32_t m = s == 0 ? 0x0000FFFFu : s == 1 ? 0x00FF00FFu : s == 2 ? 0x0F0F0F0Fu
                          : s == 3 ? 0x33333333u : 0x55555555u;
#pragma unroll
        for (int k = 0; k < 32; k++) {
            if (k & j) continue;
            const uint32_t t = ((a[k] >> j) ^ a[k + j]) & m;
            a[k + j] ^= t;
            a[k] ^= t << j;
        }
    }
}

// Same transpose with the two byte-granular stages done by PRMT and the sub-byte stages as
// shift + LOP3 pairs (the masks satisfy m << j == ~m): 256 instructions instead of ~400.
// (a & M) | (b & ~M) as ONE LOP3 with the mask as its immediate.  Written as C, nvcc emits two (a LOP3 takes one
// immediate and it materialises M and ~M separately): 192 instead of 96 LOP3 per 32x32 transpose, on a pipe that
// issues one warp instruction every other cycle.
template <uint32_t M>
__device__ __forceinline__ uint32_t bit_select(uint32_t a, uint32_t b)
{
    uint32_t d;
    asm("lop3.b32 %0, %1, %2, %3, 0xE4;" : "=r"(d) : "r"(a), "r"(b), "n"(M));
    return d;
}

// x >> J as a multiply-high by 2^(32-J) read from constant memory: IMAD.HI on the FMA pipe instead of SHF on the ALU
// pipe, which carries everything else of a transpose (the multiplier must not be a literal, or ptxas turns the
// multiply back into the shift).  Index: J = 4, 2, 1 -> 0, 1, 2.
static __constant__ uint32_t bbb_shr_mul[3] = {1u << 28, 1u << 30, 1u << 31};

template <int J, uint32_t M, bool FMA_SHR, int N>
__device__ __forceinline__ void transpose_bits_stage(uint32_t (&a)[N])
{
    const uint32_t mul = FMA_SHR ? bbb_shr_mul[J == 4 ? 0 : J == 2 ? 1 : 2] : 0u;
#pragma unroll
    for (int k = 0; k < N; k++) {
        if (k & J) continue;
        const uint32_t x = a[k], y = a[k + J];
        a[k] = bit_select<M>(x, y << J);
        a[k + J] = bit_select<M>(FMA_SHR ? __umulhi(x, mul) : (x >> J), y);
    }
}

template <bool FMA_SHR = false>
__device__ __forceinline__ void transpose32_fast(uint32_t (&a)[32])
{
#pragma unroll
    for (int k = 0; k < 16; k++) {
        const uint32_t lo = __byte_perm(a[k], a[k + 16], 0x5410), hi = __byte_perm(a[k], a[k + 16], 0x7632);
        a[k] = lo;
        a[k + 16] = hi;
    }
#pragma unroll
    for (int k = 0; k < 32; k++) {
        if (k & 8) continue;
        const uint32_t lo = __byte_perm(a[k], a[k + 8], 0x6240), hi = __byte_perm(a[k], a[k + 8], 0x7351);
        a[k] = lo;
        a[k + 8] = hi;
    }
    transpose_bits_stage<4, 0x0F0F0F0Fu, FMA_SHR>(a);
    transpose_bits_stage<2, 0x33333333u, FMA_SHR>(a);
    transpose_bits_stage<1, 0x55555555u, FMA_SHR>(a);
}

// The 32x32 transpose in two halves that commute: the three sub-byte stages on all 32 rows (here), and the two
// byte-granular stages, which only mix rows k, k+8, k+16, k+24 (byte4x4_transpose on those four).
template <bool WITH_STAGE1 = true>
__device__ __forceinline__ void transpose32_bit_stages(uint32_t (&a)[32])
{
    transpose_bits_stage<4, 0x0F0F0F0Fu, false>(a);
    transpose_bits_stage<2, 0x33333333u, false>(a);
    if (WITH_STAGE1) transpose_bits_stage<1, 0x55555555u, false>(a);
}

// the last sub-byte stage on one pair of rows (k even, k + 1)
__device__ __forceinline__ void bit_stage1_pair(uint32_t &x, uint32_t &y)
{
    const uint32_t a = bit_select<0x55555555u>(x, y << 1), b = bit_select<0x55555555u>(x >> 1, y);
    x = a;
    y = b;
}

__device__ __forceinline__ void byte4x4_transpose(uint32_t &r0, uint32_t &r1, uint32_t &r2, uint32_t &r3)
{
    const uint32_t lo02 = __byte_perm(r0, r2, 0x5410), hi02 = __byte_perm(r0, r2, 0x7632);
    const uint32_t lo13 = __byte_perm(r1, r3, 0x5410), hi13 = __byte_perm(r1, r3, 0x7632);
    r0 = __byte_perm(lo02, lo13, 0x6240);
    r1 = __byte_perm(lo02, lo13, 0x7351);
    r2 = __byte_perm(hi02, hi13, 0x6240);
    r3 = __byte_perm(hi02, hi13, 0x7351);
}

// acc[0..n] += hist[0..n]; acc[n+1] += sum g*h; acc[n+2] += sum g^2*h with g = bin - n/2.
// Exact integer arithmetic: the moments of the sample set are functions of its histogram.
static __global__ void fold_hist_kernel(const unsigned long long *__restrict__ hist, uint32_t n,
                                 unsigned long long *__restrict__ acc)
{
    __shared__ long long s1[512];
    __shared__ unsigned long long s2[512];
    const uint32_t tid = threadIdx.x;
    long long a = 0;
    unsigned long long b = 0;
    for (uint32_t i = tid; i <= n; i += blockDim.x) {
        const unsigned long long h = hist[i];
        const long long g = (long long)i - (long long)(n / 2);
        acc[i] += h;
        a += g * (long long)h;
        b += (unsigned long long)(g * g) * h;
    }
    s1[tid] = a;
    s2[tid] = b;
    __syncthreads();
    for (uint32_t st = blockDim.x / 2; st; st >>= 1) {
        if (tid < st) { s1[tid] += s1[tid + st]; s2[tid] += s2[tid + st]; }
        __syncthreads();
    }
    if (tid == 0) {
        acc[n + 1] += (unsigned long long)s1[0];
        acc[n + 2] += s2[0];
    }
}
