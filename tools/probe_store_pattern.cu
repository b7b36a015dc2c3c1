// Memory-side ceiling of the raw-word stream's store pattern, without any of its arithmetic.
//
// clt256_words2_kernel (basebandboard_b200/csrc/clt_spec.cu) keeps 148 x 128 x 32 = 606 208 sub-streams open and
// appends one 32-byte clock to each of them per iteration.  This probe issues exactly those stores (same grid, same
// addresses, same lane-pair arrangement and cache policy) with constant data, and then varies one thing at a time:
//   visit   bytes appended to a stream per visit (32 = words2; 64 / 128 / 256 = what a K-clock stage would give)
//   spt     streams per thread (32 = words2; fewer = fewer open streams at the same grid)
//   policy  0 st.cs, 1 default, 2 L2 evict_last
// build:  nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o build/probe_store_pattern tools/probe_store_pattern.cu
// run:    build/probe_store_pattern            (prints one line per combination; 8 GiB written per launch)
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <cstdlib>

#define CHECK(x)                                                                          \
    do {                                                                                  \
        cudaError_t e_ = (x);                                                             \
        if (e_ != cudaSuccess) {                                                          \
            fprintf(stderr, "%s:%d %s\n", __FILE__, __LINE__, cudaGetErrorString(e_));    \
            exit(1);                                                                      \
        }                                                                                 \
    } while (0)

template <int POLICY>
__device__ __forceinline__ void store16(uint4 *p, uint4 v, uint64_t pol)
{
    if (POLICY == 0) __stcs(p, v);
    else if (POLICY == 1) *p = v;
    else
        asm volatile("st.global.L2::cache_hint.v4.u32 [%0], {%1, %2, %3, %4}, %5;" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z),
                     "r"(v.w), "l"(pol) : "memory");
}

// One warp serves the 32 * SPT streams of its 32 "owner" threads.  A visit of VISIT bytes is written by G = VISIT / 16
// adjacent lanes (16 bytes each), so one warp store covers 32 / G visits.  Streams of owner thread u: (t*SPT + j),
// each `visits * VISIT` bytes long.
template <int POLICY, int VISIT, int SPT>
__global__ void __launch_bounds__(128, 1) store_kernel(uint4 *out, uint32_t visits)
{
    constexpr int G = VISIT / 16;            // lanes per visit
    constexpr int VPI = 32 / G;              // visits per warp instruction
    uint64_t pol = 0;
    if (POLICY == 2) asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(pol));
    const uint32_t lane = threadIdx.x & 31u, q = lane / G, h = lane % G;
    const uint32_t wbase = blockIdx.x * 128u + (threadIdx.x & ~31u);
    const size_t stream_u4 = (size_t)visits * G;                         // uint4 per stream
    const uint4 v = make_uint4(lane, blockIdx.x, threadIdx.x, 0x5a5a5a5au);
#pragma unroll 1
    for (uint32_t it = 0; it < visits; it++) {
#pragma unroll 1
        for (uint32_t u = q; u < 32u; u += VPI) {                        // owner thread within the warp
            uint4 *base = out + (size_t)(wbase + u) * SPT * stream_u4 + (size_t)it * G + h;
#pragma unroll
            for (int j = 0; j < SPT; j++) store16<POLICY>(base + (size_t)j * stream_u4, v, pol);
        }
    }
}

template <int POLICY, int VISIT, int SPT>
static void run(uint4 *buf, size_t bytes, int blocks)
{
    const size_t streams = (size_t)blocks * 128 * SPT;
    const uint32_t visits = (uint32_t)(bytes / streams / VISIT);
    const size_t written = streams * visits * VISIT;
    cudaEvent_t a, b;
    CHECK(cudaEventCreate(&a));
    CHECK(cudaEventCreate(&b));
    float best = 1e30f;
    for (int rep = 0; rep < 3; rep++) {
        CHECK(cudaEventRecord(a));
        store_kernel<POLICY, VISIT, SPT><<<blocks, 128>>>(buf, visits);
        CHECK(cudaEventRecord(b));
        CHECK(cudaEventSynchronize(b));
        float ms;
        CHECK(cudaEventElapsedTime(&ms, a, b));
        if (ms < best) best = ms;
    }
    CHECK(cudaGetLastError());
    printf("{\"policy\": %d, \"visit_bytes\": %d, \"streams_per_thread\": %d, \"open_streams\": %zu, \"stream_bytes\": %zu, "
           "\"GBps\": %.1f}\n",
           POLICY, VISIT, SPT, streams, (size_t)visits * VISIT, written / (best * 1e-3) / 1e9);
    fflush(stdout);
}

template <int POLICY>
static void sweep(uint4 *buf, size_t bytes, int blocks)
{
    run<POLICY, 32, 32>(buf, bytes, blocks);
    run<POLICY, 64, 32>(buf, bytes, blocks);
    run<POLICY, 128, 32>(buf, bytes, blocks);
    run<POLICY, 256, 32>(buf, bytes, blocks);
    run<POLICY, 32, 16>(buf, bytes, blocks);
    run<POLICY, 32, 8>(buf, bytes, blocks);
    run<POLICY, 32, 4>(buf, bytes, blocks);
    run<POLICY, 128, 8>(buf, bytes, blocks);
}

int main(int argc, char **argv)
{
    const size_t bytes = (size_t)8 << 30;
    int dev = 0, sms = 0;
    CHECK(cudaSetDevice(dev));
    CHECK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    const int blocks = argc > 1 ? atoi(argv[1]) : sms;
    uint4 *buf;
    CHECK(cudaMalloc(&buf, bytes));
    CHECK(cudaMemset(buf, 0, bytes));
    sweep<2>(buf, bytes, blocks);
    sweep<1>(buf, bytes, blocks);
    sweep<0>(buf, bytes, blocks);
    // two CTAs' worth of streams per SM (words4 / words5 geometry)
    run<2, 32, 32>(buf, bytes, 2 * blocks);
    run<2, 128, 32>(buf, bytes, 2 * blocks);
    CHECK(cudaFree(buf));
    return 0;
}
