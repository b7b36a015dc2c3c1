// bbb_internal.h -- shared between the translation units of libbbb_b200.so.
#pragma once

#include <cuda_runtime.h>

#include <cstdarg>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <functional>
#include <string>
#include <vector>

#include "../../include/bbb_b200.h"

#define BBB_MAX_N 512
#define BBB_MAX_NW64 (BBB_MAX_N / 64)
#define BBB_MAX_NW32 (BBB_MAX_N / 32)

// Dense n x n GF(2) matrix on the host, row-major, 64-bit words, bit c of row r
// = word c/64 bit c%64.  Used only to precompute jump-ahead operators A^e
// ("host precomputes A^(2^i) by repeated squaring").
struct HostMat {
    uint32_t n = 0, nw = 0;
    std::vector<uint64_t> w;  // n * nw
    void init(uint32_t n_) { n = n_; nw = (n_ + 63) / 64; w.assign((size_t)n * nw, 0); }
    uint64_t *row(uint32_t r) { return &w[(size_t)r * nw]; }
    const uint64_t *row(uint32_t r) const { return &w[(size_t)r * nw]; }
    void set_identity();
    static void mul(HostMat &c, const HostMat &a, const HostMat &b);   // c = a.b
    void apply(const uint64_t *x, uint64_t *y) const;                  // y = M.x (may alias)
    // Column-major 32-bit words for the seeding kernel: out[c*nw32 + w].
    void to_colmajor32(std::vector<uint32_t> &out) const;
};

struct bbb_ctx {
    int device = 0;
    cudaStream_t stream = nullptr;
    cudaStream_t stream2 = nullptr;    // second queue: overlapped batches (rnghunt) and chunked D2H copies (streams)
    cudaStream_t stream3 = nullptr;    // third queue: the ragged tail of a CLT / word-stream call beside its main launch
    cudaEvent_t ev_tail_fork = nullptr, ev_tail_join = nullptr;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr, evt0 = nullptr, evt1 = nullptr;
    cudaEvent_t ev_fork = nullptr, ev_join = nullptr, ev_slot[2] = {nullptr, nullptr}, ev_aux[2] = {nullptr, nullptr};
    uint64_t h2d_bytes = 0, d2h_bytes = 0;
    int sm_count = 0;
    std::string err;
    double last_ms = 0.0;
    uint32_t last_launches = 0;
    const char *last_kernel = "none";

    // loaded LUTOPT recurrence
    uint32_t n = 0;
    std::vector<uint32_t> row_ptr;
    std::vector<uint16_t> taps;
    int shipped = 0;              // n of the shipped matrix this equals, else 0
    // tuning knobs (bbb_set_tuning): tests use them to force a code path
    int tune_no_spec = 0;               // 1: never use the specialised n256 kernel
    int tune_spec_variant = 2;          // fused n256 kernel: 1 two-pass (clt256_gen.inc), 2 one-pass (clt256_v2_gen.inc)
    uint32_t tune_spec_min_steps = 512; // fewest clocks per stream worth a specialised launch
    uint32_t tune_spec_max_blocks = 0;  // 0: one CTA per SM
    uint64_t tune_slicer_chunk_bits = 0; // 0: 2^27 PRBS bits per pass of the slicer-loop composite
    int tune_words_variant = 0;         // raw-word kernel: 0 default, 2 clt256_words2_kernel, 4 clt256_words4_kernel (two CTAs per SM)
    int tune_words_b192 = 0;            // 1: raw-word stream on the 192-thread experiment (clt256_words3_kernel)
    int tune_seed_single = 0;           // 1: always the one-group-per-warp seeding kernel
    int tune_dec_full_stream = 0;       // 1: decimated samples by the full stream + pick (never clt256_dec_kernel)
    int nest = 0;                       // > 0 inside a composite ABI call (begin_call/end_call are the outer call's)
    uint32_t tune_prbs_chunk_words = 0; // 0: automatic detector chunk length (64-bit words)
    uint32_t tune_prbs_stream_reps = 0; // 0: automatic stream-kernel repetitions per warp
    int tune_hunt_smem_taps = 0;        // 1: never use the taps-in-registers front kernel
    int tune_hunt_slow_check = 0;       // 1: always use the shift-and-add primitivity kernel
    uint32_t tune_hunt_batch = 0;       // 0: 2^20 candidates per rnghunt batch
    uint64_t tune_stream_chunk = 0;     // bytes per chunk of the host-buffer stream calls (0: 64 MiB / 256 MiB)
    uint32_t tune_host_copy_threads = 0; // threads copying a chunk out of the pinned staging buffer (0: up to 8)
    uint32_t tune_hunt_pow_k = 0;       // clocks per pass of hunt_front_pow_kernel: 0/8, or 4
    uint32_t tune_hunt_cpw = 0;         // candidates per warp pass of hunt_front_warp_kernel: 0/32, or 16
    uint32_t last_rounds = 0;           // kernel rounds of the last bbb_prbs_detect
    // PRBS jump tables per k (prbs.cu), built on first use
    void *prbs_host[8] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
    void *prbs_dev[8] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
    // cached jump-ahead operators on the device (scratch slot 0)
    struct JumpSet {              // A^(L*2^b), b < nbits, column-major u32, on the device
        uint64_t L = 0;
        uint32_t nbits = 0;
        uint32_t *dev = nullptr;
        size_t bytes = 0;
        uint64_t stamp = 0;
    } jump_sets[24];               // a chunked host-buffer stream call repeats the same handful of geometries chunk after chunk
    uint64_t jump_stamp = 0;
    HostMat A;                    // dense A
    std::vector<HostMat> Apow2;   // A^(2^i), grown on demand
    uint32_t *d_row_ptr = nullptr;   // device copies for the generic kernel
    uint16_t *d_taps = nullptr;

    // scratch device buffers (grown on demand)
    void *d_scratch[16] = {};
    size_t scratch_bytes[16] = {};

    int fail(int code, const char *fmt, ...);
    int cuda_fail(cudaError_t e, const char *what);
    void *scratch(int slot, size_t bytes);   // nullptr on failure (err set)
    // pinned (page-locked, device-visible) host buffers, grown on demand
    void *h_pinned[4] = {};
    size_t pinned_bytes[4] = {};
    void *pinned(int slot, size_t bytes);    // nullptr on failure (err set)
    // counted copies on the ctx's stream
    cudaError_t h2d(void *dst, const void *src, size_t bytes)
    {
        h2d_bytes += bytes;
        return cudaMemcpyAsync(dst, src, bytes, cudaMemcpyHostToDevice, stream);
    }
    cudaError_t d2h(void *dst, const void *src, size_t bytes)
    {
        d2h_bytes += bytes;
        return cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToHost, stream);
    }
    void begin_call();                       // resets timing counters, records ev0
    void end_call();                         // records ev1 (call bbb_sync before reading ms)
    const HostMat &pow2(uint32_t i);         // A^(2^i)
    void power(uint64_t e, HostMat &out);    // A^e
};

#define BBB_CUDA(ctx, call)                                       \
    do {                                                          \
        cudaError_t e__ = (call);                                 \
        if (e__ != cudaSuccess) return (ctx)->cuda_fail(e__, #call); \
    } while (0)

// api.cu: a materialised stream delivered to caller-owned (pageable) host memory.  `gen(off, bytes, dev)` queues
// the kernels that write bytes [off, off+bytes) of the stream to `dev` on ctx->stream; chunks alternate between
// two device buffers (scratch slot `dev_slot`) and two pinned staging buffers, the D2H copy of chunk i runs on
// ctx->stream2 beside the kernels of chunk i+1, and the host thread(s) copy chunk i-1 out of the staging buffer
// meanwhile.  One timing region (bbb_last_device_ms) spans the call.  `align`: chunk boundaries in bytes.
int bbb_stream_to_host(bbb_ctx *ctx, int dev_slot, void *out, size_t total_bytes, size_t align, size_t chunk_default,
                       const std::function<int(size_t off, size_t bytes, void *dev)> &gen);

// prbs.cu
void bbb_prbs_free_tables(bbb_ctx *ctx);

// clt.cu
int bbb_clt_run_device(bbb_ctx *ctx, const uint64_t *x0, uint64_t first, uint64_t count, uint32_t model,
                       uint64_t *d_hist /*n+1 u64 or null*/, int8_t *d_out8, int16_t *d_out16,
                       uint32_t noise_var, uint32_t *d_out_words32);
// d_out[j] = g[first + j*spb + delay], j < nd (int8, gateware widths)
int bbb_clt_run_decimated(bbb_ctx *ctx, const uint64_t *x0, uint64_t first, uint64_t nd, uint32_t spb, uint32_t delay,
                          int8_t *d_out);
