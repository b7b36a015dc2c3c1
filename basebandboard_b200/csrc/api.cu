// api.cu -- context management, host-side GF(2) jump-ahead operators, LUTOPT
// loading, derived statistics and the LOP3 roofline micro-benchmark.
#include <cmath>
#include <thread>

#include "bbb_internal.h"
#include "recurrences_data.inc"

// ---------------------------------------------------------------------------
// HostMat
// ---------------------------------------------------------------------------
void HostMat::set_identity()
{
    std::fill(w.begin(), w.end(), 0);
    for (uint32_t r = 0; r < n; r++) row(r)[r >> 6] |= 1ull << (r & 63);
}

// Row r of a.b is the XOR of the rows j of b selected by the set bits of a[r].
void HostMat::mul(HostMat &c, const HostMat &a, const HostMat &b)
{
    HostMat out;
    out.init(a.n);
    const uint32_t n = a.n, nw = a.nw;
    for (uint32_t r = 0; r < n; r++) {
        uint64_t acc[BBB_MAX_NW64] = {0};
        const uint64_t *ar = a.row(r);
        for (uint32_t wi = 0; wi < nw; wi++) {
            uint64_t bits = ar[wi];
            while (bits) {
                uint32_t j = wi * 64 + (uint32_t)__builtin_ctzll(bits);
                bits &= bits - 1;
                const uint64_t *br = b.row(j);
                for (uint32_t k = 0; k < nw; k++) acc[k] ^= br[k];
            }
        }
        memcpy(out.row(r), acc, nw * sizeof(uint64_t));
    }
    c = std::move(out);
}

void HostMat::apply(const uint64_t *x, uint64_t *y) const
{
    uint64_t out[BBB_MAX_NW64] = {0};
    for (uint32_t r = 0; r < n; r++) {
        uint64_t p = 0;
        const uint64_t *rr = row(r);
        for (uint32_t k = 0; k < nw; k++) p ^= rr[k] & x[k];
        out[r >> 6] |= (uint64_t)(__builtin_popcountll(p) & 1) << (r & 63);
    }
    memcpy(y, out, nw * sizeof(uint64_t));
}

void HostMat::to_colmajor32(std::vector<uint32_t> &out) const
{
    const uint32_t nw32 = (n + 31) / 32;
    out.assign((size_t)n * nw32, 0);
    for (uint32_t r = 0; r < n; r++)
        for (uint32_t c = 0; c < n; c++)
            if ((row(r)[c >> 6] >> (c & 63)) & 1) out[(size_t)c * nw32 + (r >> 5)] |= 1u << (r & 31);
}

// ---------------------------------------------------------------------------
// ctx
// ---------------------------------------------------------------------------
int bbb_ctx::fail(int code, const char *fmt, ...)
{
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    err = buf;
    return code;
}

int bbb_ctx::cuda_fail(cudaError_t e, const char *what)
{
    return fail(BBB_ECUDA, "CUDA error %d (%s) in %s", (int)e, cudaGetErrorString(e), what);
}

void *bbb_ctx::scratch(int slot, size_t bytes)
{
    if (bytes <= scratch_bytes[slot] && d_scratch[slot]) return d_scratch[slot];
    if (d_scratch[slot]) cudaFree(d_scratch[slot]);
    d_scratch[slot] = nullptr;
    scratch_bytes[slot] = 0;
    size_t want = bytes < 4096 ? 4096 : bytes;
    cudaError_t e = cudaMalloc(&d_scratch[slot], want);
    if (e != cudaSuccess) {
        cuda_fail(e, "cudaMalloc(scratch)");
        d_scratch[slot] = nullptr;
        return nullptr;
    }
    scratch_bytes[slot] = want;
    return d_scratch[slot];
}

void *bbb_ctx::pinned(int slot, size_t bytes)
{
    if (bytes <= pinned_bytes[slot] && h_pinned[slot]) return h_pinned[slot];
    if (h_pinned[slot]) cudaFreeHost(h_pinned[slot]);
    h_pinned[slot] = nullptr;
    pinned_bytes[slot] = 0;
    size_t want = bytes < 4096 ? 4096 : bytes;
    cudaError_t e = cudaHostAlloc(&h_pinned[slot], want, cudaHostAllocPortable | cudaHostAllocMapped);
    if (e != cudaSuccess) {
        cuda_fail(e, "cudaHostAlloc(pinned)");
        h_pinned[slot] = nullptr;
        return nullptr;
    }
    pinned_bytes[slot] = want;
    return h_pinned[slot];
}

void bbb_ctx::begin_call()
{
    if (nest) return;      // inside a composite call: one timing region, counters keep adding up
    last_ms = 0.0;
    last_launches = 0;
    h2d_bytes = d2h_bytes = 0;
    cudaEventRecord(ev0, stream);
}

void bbb_ctx::end_call()
{
    if (nest) return;
    cudaEventRecord(ev1, stream);
}

const HostMat &bbb_ctx::pow2(uint32_t i)
{
    if (Apow2.empty()) Apow2.push_back(A);
    while (Apow2.size() <= i) {
        HostMat sq;
        HostMat::mul(sq, Apow2.back(), Apow2.back());
        Apow2.push_back(std::move(sq));
    }
    return Apow2[i];
}

void bbb_ctx::power(uint64_t e, HostMat &out)
{
    HostMat acc;
    acc.init(n);
    acc.set_identity();
    bool ident = true;
    for (uint32_t i = 0; e; i++, e >>= 1) {
        if (!(e & 1)) continue;
        if (ident) { acc = pow2(i); ident = false; }
        else HostMat::mul(acc, pow2(i), acc);
    }
    out = std::move(acc);
}

// ---------------------------------------------------------------------------
// chunked, double-buffered delivery of a materialised stream to pageable host memory
// ---------------------------------------------------------------------------
static void host_copy_parallel(char *dst, const char *src, size_t bytes, uint32_t threads)
{
    if (threads <= 1 || bytes < (8u << 20)) { memcpy(dst, src, bytes); return; }
    const size_t per = ((bytes + threads - 1) / threads + 4095) & ~(size_t)4095;
    std::vector<std::thread> pool;
    for (uint32_t t = 1; t < threads; t++) {
        const size_t lo = (size_t)t * per;
        if (lo >= bytes) break;
        const size_t len = bytes - lo < per ? bytes - lo : per;
        pool.emplace_back([=] { memcpy(dst + lo, src + lo, len); });
    }
    memcpy(dst, src, bytes < per ? bytes : per);
    for (auto &th : pool) th.join();
}

int bbb_stream_to_host(bbb_ctx *c, int dev_slot, void *out, size_t total, size_t align, size_t chunk_default,
                       const std::function<int(size_t off, size_t bytes, void *dev)> &gen)
{
    BBB_CUDA(c, cudaSetDevice(c->device));
    if (!align) align = 1;
    size_t chunk = c->tune_stream_chunk ? (size_t)c->tune_stream_chunk : chunk_default;
    chunk = chunk / align * align;
    if (chunk < align) chunk = align;
    if (chunk > total) chunk = (total + align - 1) / align * align;
    if (!chunk) chunk = align;
    const size_t slot_bytes = (chunk + 255) & ~(size_t)255;
    uint32_t threads = c->tune_host_copy_threads;
    if (!threads) {
        threads = std::thread::hardware_concurrency();
        threads = threads > 8 ? 8 : (threads ? threads : 1);
    }
    // a caller that hands in page-locked memory (cudaHostAlloc / cudaHostRegister, e.g. a pinned torch tensor) gets the
    // chunks copied straight into it: no staging buffer, no host-side copy
    bool direct = false;
    {
        cudaPointerAttributes attr;
        if (total && cudaPointerGetAttributes(&attr, out) == cudaSuccess) direct = attr.type == cudaMemoryTypeHost;
        cudaGetLastError();                          // an unregistered pointer is not an error here
    }
    c->begin_call();
    struct Nest { bbb_ctx *c; Nest(bbb_ctx *cc) : c(cc) { c->nest++; } ~Nest() { c->nest--; } } nest(c);
    char *d = (char *)c->scratch(dev_slot, 2 * slot_bytes);
    char *h = direct ? (char *)out : (char *)c->pinned(1, 2 * slot_bytes);
    if (!d || !h) return BBB_ECUDA;
    const size_t nchunks = (total + chunk - 1) / chunk;
    auto drain = [&](size_t i) -> int {
        const int slot = (int)(i & 1);
        const size_t off = i * chunk, bytes = total - off < chunk ? total - off : chunk;
        BBB_CUDA(c, cudaEventSynchronize(c->ev_slot[slot]));      // (direct: the device buffer is free again)
        if (!direct) host_copy_parallel((char *)out + off, h + slot * slot_bytes, bytes, threads);
        return BBB_OK;
    };
    for (size_t i = 0; i < nchunks; i++) {
        const int slot = (int)(i & 1);
        const size_t off = i * chunk, bytes = total - off < chunk ? total - off : chunk;
        int rc = gen(off, bytes, d + slot * slot_bytes);             // kernels of chunk i on c->stream
        if (rc) return rc;
        BBB_CUDA(c, cudaEventRecord(c->ev_aux[slot], c->stream));
        BBB_CUDA(c, cudaStreamWaitEvent(c->stream2, c->ev_aux[slot], 0));
        BBB_CUDA(c, cudaMemcpyAsync(direct ? h + off : h + slot * slot_bytes, d + slot * slot_bytes, bytes,
                                    cudaMemcpyDeviceToHost, c->stream2));
        c->d2h_bytes += bytes;
        BBB_CUDA(c, cudaEventRecord(c->ev_slot[slot], c->stream2));
        if (i) { rc = drain(i - 1); if (rc) return rc; }             // chunk i-1 leaves the staging buffer meanwhile
    }
    if (nchunks) { int rc = drain(nchunks - 1); if (rc) return rc; }
    BBB_CUDA(c, cudaEventRecord(c->ev_join, c->stream2));
    BBB_CUDA(c, cudaStreamWaitEvent(c->stream, c->ev_join, 0));
    c->nest--;          // close the outer timing region (the guard's decrement is balanced below)
    c->end_call();
    c->nest++;
    BBB_CUDA(c, cudaStreamSynchronize(c->stream));
    return BBB_OK;
}

extern "C" {

int bbb_version(void) { return 200; }

int bbb_device_count(void)
{
    int count = 0;
    if (cudaGetDeviceCount(&count) != cudaSuccess) return 0;
    return count;
}

int bbb_create(int device, bbb_ctx **out)
{
    if (!out) return BBB_EINVAL;
    *out = nullptr;
    int count = 0;
    cudaError_t e = cudaGetDeviceCount(&count);
    if (e != cudaSuccess || count <= 0 || device < 0 || device >= count) return BBB_ECUDA;
    if (cudaSetDevice(device) != cudaSuccess) return BBB_ECUDA;
    bbb_ctx *c = new (std::nothrow) bbb_ctx();
    if (!c) return BBB_ENOMEM;
    c->device = device;
    cudaDeviceProp prop;
    if (cudaGetDeviceProperties(&prop, device) != cudaSuccess) { delete c; return BBB_ECUDA; }
    if (prop.major < 10) { delete c; return BBB_EUNSUPPORTED; }   // built for sm_100a only
    c->sm_count = prop.multiProcessorCount;
    if (cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking) != cudaSuccess ||
        cudaStreamCreateWithFlags(&c->stream2, cudaStreamNonBlocking) != cudaSuccess ||
        cudaStreamCreateWithFlags(&c->stream3, cudaStreamNonBlocking) != cudaSuccess ||
        cudaEventCreateWithFlags(&c->ev_tail_fork, cudaEventDisableTiming) != cudaSuccess ||
        cudaEventCreateWithFlags(&c->ev_tail_join, cudaEventDisableTiming) != cudaSuccess ||
        cudaEventCreateWithFlags(&c->ev_fork, cudaEventDisableTiming) != cudaSuccess ||
        cudaEventCreateWithFlags(&c->ev_join, cudaEventDisableTiming) != cudaSuccess ||
        cudaEventCreateWithFlags(&c->ev_slot[0], cudaEventDisableTiming) != cudaSuccess ||
        cudaEventCreateWithFlags(&c->ev_slot[1], cudaEventDisableTiming) != cudaSuccess ||
        cudaEventCreateWithFlags(&c->ev_aux[0], cudaEventDisableTiming) != cudaSuccess ||
        cudaEventCreateWithFlags(&c->ev_aux[1], cudaEventDisableTiming) != cudaSuccess ||
        cudaEventCreate(&c->ev0) != cudaSuccess || cudaEventCreate(&c->ev1) != cudaSuccess ||
        cudaEventCreate(&c->evt0) != cudaSuccess || cudaEventCreate(&c->evt1) != cudaSuccess) {
        delete c;
        return BBB_ECUDA;
    }
    *out = c;
    return BBB_OK;
}

void bbb_destroy(bbb_ctx *c)
{
    if (!c) return;
    cudaSetDevice(c->device);
    cudaStreamSynchronize(c->stream);
    if (c->stream2) cudaStreamSynchronize(c->stream2);
    if (c->stream3) cudaStreamSynchronize(c->stream3);
    if (c->ev_tail_fork) cudaEventDestroy(c->ev_tail_fork);
    if (c->ev_tail_join) cudaEventDestroy(c->ev_tail_join);
    if (c->stream3) cudaStreamDestroy(c->stream3);
    for (int i = 0; i < 16; i++) if (c->d_scratch[i]) cudaFree(c->d_scratch[i]);
    for (int i = 0; i < 4; i++) if (c->h_pinned[i]) cudaFreeHost(c->h_pinned[i]);
    if (c->ev_fork) cudaEventDestroy(c->ev_fork);
    if (c->ev_join) cudaEventDestroy(c->ev_join);
    for (int i = 0; i < 2; i++) if (c->ev_slot[i]) cudaEventDestroy(c->ev_slot[i]);
    for (int i = 0; i < 2; i++) if (c->ev_aux[i]) cudaEventDestroy(c->ev_aux[i]);
    if (c->stream2) cudaStreamDestroy(c->stream2);
    if (c->d_row_ptr) cudaFree(c->d_row_ptr);
    if (c->d_taps) cudaFree(c->d_taps);
    bbb_prbs_free_tables(c);
    for (auto &js : c->jump_sets) if (js.dev) cudaFree(js.dev);
    cudaEventDestroy(c->ev0);
    cudaEventDestroy(c->ev1);
    cudaEventDestroy(c->evt0);
    cudaEventDestroy(c->evt1);
    cudaStreamDestroy(c->stream);
    delete c;
}

const char *bbb_last_error(const bbb_ctx *c) { return c ? c->err.c_str() : "null ctx"; }

double bbb_last_device_ms(const bbb_ctx *c)
{
    if (!c) return 0.0;
    float ms = 0.f;
    if (cudaEventSynchronize(c->ev1) != cudaSuccess) return 0.0;
    if (cudaEventElapsedTime(&ms, c->ev0, c->ev1) != cudaSuccess) return 0.0;
    return (double)ms;
}

uint32_t bbb_last_launches(const bbb_ctx *c) { return c ? c->last_launches : 0; }
const char *bbb_last_kernel(const bbb_ctx *c) { return c ? c->last_kernel : "none"; }

uint32_t bbb_last_rounds(const bbb_ctx *c) { return c ? c->last_rounds : 0; }

uint64_t bbb_last_h2d_bytes(const bbb_ctx *c) { return c ? c->h2d_bytes : 0; }
uint64_t bbb_last_d2h_bytes(const bbb_ctx *c) { return c ? c->d2h_bytes : 0; }

int bbb_set_tuning(bbb_ctx *c, const char *key, int64_t value)
{
    if (!c || !key) return BBB_EINVAL;
    if (!strcmp(key, "clt_no_spec")) c->tune_no_spec = value != 0;
    else if (!strcmp(key, "clt_spec_variant")) c->tune_spec_variant = value == 1 ? 1 : 2;
    else if (!strcmp(key, "clt_spec_min_steps")) c->tune_spec_min_steps = value < 1 ? 1u : (uint32_t)value;
    else if (!strcmp(key, "clt_spec_max_blocks")) c->tune_spec_max_blocks = value < 0 ? 0u : (uint32_t)value;
    else if (!strcmp(key, "words_variant")) c->tune_words_variant = (int)value;
    else if (!strcmp(key, "words_b192")) c->tune_words_b192 = value != 0;
    else if (!strcmp(key, "seed_single")) c->tune_seed_single = value != 0;
    else if (!strcmp(key, "dec_full_stream")) c->tune_dec_full_stream = value != 0;
    else if (!strcmp(key, "slicer_chunk_bits")) c->tune_slicer_chunk_bits = value < 64 ? 0ull : ((uint64_t)value & ~63ull);
    else if (!strcmp(key, "prbs_chunk_words")) c->tune_prbs_chunk_words = value < 0 ? 0u : (uint32_t)value;
    else if (!strcmp(key, "hunt_smem_taps")) c->tune_hunt_smem_taps = (int)value;
    else if (!strcmp(key, "hunt_slow_check")) c->tune_hunt_slow_check = value != 0;
    else if (!strcmp(key, "hunt_batch")) c->tune_hunt_batch = value < 0 ? 0u : (uint32_t)value;
    else if (!strcmp(key, "stream_chunk_bytes")) c->tune_stream_chunk = value < 0 ? 0ull : (uint64_t)value;
    else if (!strcmp(key, "host_copy_threads")) c->tune_host_copy_threads = value < 0 ? 0u : (uint32_t)value;
    else if (!strcmp(key, "hunt_pow_k")) c->tune_hunt_pow_k = value == 4 ? 4u : 0u;
    else if (!strcmp(key, "hunt_cpw")) c->tune_hunt_cpw = value == 16 ? 16u : 0u;
    else if (!strcmp(key, "prbs_stream_reps")) c->tune_prbs_stream_reps = value < 0 ? 0u : (uint32_t)value;
    else return c->fail(BBB_EINVAL, "unknown tuning key %s", key);
    return BBB_OK;
}

int bbb_timer_start(bbb_ctx *c)
{
    if (!c) return BBB_EINVAL;
    BBB_CUDA(c, cudaSetDevice(c->device));
    BBB_CUDA(c, cudaEventRecord(c->evt0, c->stream));
    return BBB_OK;
}

int bbb_timer_stop(bbb_ctx *c, double *ms)
{
    if (!c || !ms) return BBB_EINVAL;
    BBB_CUDA(c, cudaEventRecord(c->evt1, c->stream));
    BBB_CUDA(c, cudaEventSynchronize(c->evt1));
    float f = 0.f;
    BBB_CUDA(c, cudaEventElapsedTime(&f, c->evt0, c->evt1));
    *ms = (double)f;
    return BBB_OK;
}

int bbb_sync(bbb_ctx *c)
{
    if (!c) return BBB_EINVAL;
    BBB_CUDA(c, cudaStreamSynchronize(c->stream));
    return BBB_OK;
}

// ---------------------------------------------------------------------------
// LUTOPT
// ---------------------------------------------------------------------------
int bbb_lutopt_load(bbb_ctx *c, uint32_t n, const uint32_t *row_ptr, const uint16_t *taps)
{
    if (!c || !row_ptr || !taps) return BBB_EINVAL;
    if (n < 8 || n > BBB_MAX_N || (n % 8)) return c->fail(BBB_EINVAL, "n=%u: need 8 <= n <= 512, n %% 8 == 0", n);
    if (row_ptr[0] != 0) return c->fail(BBB_EINVAL, "row_ptr[0] must be 0");
    for (uint32_t r = 0; r < n; r++)
        if (row_ptr[r + 1] < row_ptr[r] || row_ptr[r + 1] - row_ptr[r] > 64)
            return c->fail(BBB_EINVAL, "row %u: bad row_ptr (row weight must be 0..64)", r);
    const uint32_t nnz = row_ptr[n];
    for (uint32_t j = 0; j < nnz; j++)
        if (taps[j] >= n) return c->fail(BBB_EINVAL, "tap %u out of range", j);
    BBB_CUDA(c, cudaSetDevice(c->device));
    c->n = n;
    c->row_ptr.assign(row_ptr, row_ptr + n + 1);
    c->taps.assign(taps, taps + nnz);
    c->A.init(n);
    for (uint32_t r = 0; r < n; r++)
        for (uint32_t j = row_ptr[r]; j < row_ptr[r + 1]; j++)
            c->A.row(r)[taps[j] >> 6] ^= 1ull << (taps[j] & 63);   // XOR: a repeated tap cancels
    c->Apow2.clear();
    for (auto &js : c->jump_sets) { js.L = 0; js.nbits = 0; }
    if (c->d_row_ptr) cudaFree(c->d_row_ptr);
    if (c->d_taps) cudaFree(c->d_taps);
    c->d_row_ptr = nullptr;
    c->d_taps = nullptr;
    BBB_CUDA(c, cudaMalloc(&c->d_row_ptr, (n + 1) * sizeof(uint32_t)));
    BBB_CUDA(c, cudaMalloc(&c->d_taps, (nnz ? nnz : 1) * sizeof(uint16_t)));
    BBB_CUDA(c, cudaMemcpy(c->d_row_ptr, row_ptr, (n + 1) * sizeof(uint32_t), cudaMemcpyHostToDevice));
    BBB_CUDA(c, cudaMemcpy(c->d_taps, taps, nnz * sizeof(uint16_t), cudaMemcpyHostToDevice));
    // is it one of the shipped matrices (specialised kernels exist for those)?
    c->shipped = 0;
    const uint32_t *sp = nullptr;
    const uint16_t *st = nullptr;
    switch (n) {
    case 16: sp = BBB_REC16_PTR; st = BBB_REC16_TAPS; break;
    case 32: sp = BBB_REC32_PTR; st = BBB_REC32_TAPS; break;
    case 64: sp = BBB_REC64_PTR; st = BBB_REC64_TAPS; break;
    case 128: sp = BBB_REC128_PTR; st = BBB_REC128_TAPS; break;
    case 192: sp = BBB_REC192_PTR; st = BBB_REC192_TAPS; break;
    case 256: sp = BBB_REC256_PTR; st = BBB_REC256_TAPS; break;
    case 512: sp = BBB_REC512_PTR; st = BBB_REC512_TAPS; break;
    default: break;
    }
    if (sp && memcmp(sp, row_ptr, (n + 1) * sizeof(uint32_t)) == 0 &&
        memcmp(st, taps, nnz * sizeof(uint16_t)) == 0)
        c->shipped = (int)n;
    return BBB_OK;
}

int bbb_lutopt_load_shipped(bbb_ctx *c, uint32_t n)
{
    if (!c) return BBB_EINVAL;
    switch (n) {
    case 16: return bbb_lutopt_load(c, n, BBB_REC16_PTR, BBB_REC16_TAPS);
    case 32: return bbb_lutopt_load(c, n, BBB_REC32_PTR, BBB_REC32_TAPS);
    case 64: return bbb_lutopt_load(c, n, BBB_REC64_PTR, BBB_REC64_TAPS);
    case 128: return bbb_lutopt_load(c, n, BBB_REC128_PTR, BBB_REC128_TAPS);
    case 192: return bbb_lutopt_load(c, n, BBB_REC192_PTR, BBB_REC192_TAPS);
    case 256: return bbb_lutopt_load(c, n, BBB_REC256_PTR, BBB_REC256_TAPS);
    case 512: return bbb_lutopt_load(c, n, BBB_REC512_PTR, BBB_REC512_TAPS);
    default: return c->fail(BBB_EINVAL, "no shipped recurrence for n=%u", n);
    }
}

int bbb_lutopt_get(const bbb_ctx *c, uint32_t *n, uint32_t *nnz, uint32_t *row_ptr, uint16_t *taps)
{
    if (!c) return BBB_EINVAL;
    if (!c->n) return BBB_ESTATE;
    if (n) *n = c->n;
    if (nnz) *nnz = (uint32_t)c->taps.size();
    if (row_ptr) memcpy(row_ptr, c->row_ptr.data(), c->row_ptr.size() * sizeof(uint32_t));
    if (taps) memcpy(taps, c->taps.data(), c->taps.size() * sizeof(uint16_t));
    return BBB_OK;
}

int bbb_lutopt_jump(bbb_ctx *c, const uint64_t *x_in, uint64_t e, uint64_t *x_out)
{
    if (!c || !x_in || !x_out) return BBB_EINVAL;
    if (!c->n) return c->fail(BBB_ESTATE, "no recurrence loaded");
    uint64_t x[BBB_MAX_NW64] = {0};
    memcpy(x, x_in, c->A.nw * sizeof(uint64_t));
    for (uint32_t i = 0; e; i++, e >>= 1)
        if (e & 1) c->pow2(i).apply(x, x);
    memcpy(x_out, x, c->A.nw * sizeof(uint64_t));
    return BBB_OK;
}

// CLTGRNG of one URNG word (rng.py:96-108): the alternating-difference tree = sum_i (-1)^popcount(i) x[i].
// Host arithmetic on one state: what the hardware shows logn clocks after reset (clt(x0), rng.py:189-190), which
// no sample index of the device stream reaches (g[i] = clt(A^(i+1) x0)).
int bbb_clt_value(uint32_t n, const uint64_t *x, uint32_t model, int32_t *g)
{
    if (!x || !g || n < 2 || n > BBB_MAX_N || (n & (n - 1)) || model > 1) return BBB_EINVAL;
    int32_t acc = 0;
    for (uint32_t i = 0; i < n; i++)
        if ((x[i >> 6] >> (i & 63)) & 1ull) acc += (__builtin_popcount(i) & 1) ? -1 : 1;
    if (model == BBB_MODEL_GATEWARE && acc == (int32_t)(n / 2)) acc = -(int32_t)(n / 2);   // Signal((logn, True)), rng.py:78
    *g = acc;
    return BBB_OK;
}

// ---------------------------------------------------------------------------
// clt-grng-evaluate.py:18-21 style statistics from the integer histogram
// ---------------------------------------------------------------------------
int bbb_clt_stats(uint32_t n, uint32_t model, const uint64_t *hist, double *s)
{
    if (!hist || !s || n < 2 || n > BBB_MAX_N || (n & (n - 1))) return BBB_EINVAL;
    const uint32_t bins = n + (model == BBB_MODEL_NUMPY ? 1 : 0);
    const double half = (double)(n / 2);
    // exact integer moments first (128-bit), then one division each
    unsigned __int128 N = 0, s2 = 0, s4 = 0;
    __int128 s1 = 0, s3 = 0;
    for (uint32_t b = 0; b < bins; b++) {
        __int128 g = (__int128)b - (__int128)(n / 2);
        unsigned __int128 h = hist[b];
        N += h;
        s1 += g * (__int128)h;
        s2 += (unsigned __int128)(g * g) * h;
        s3 += g * g * g * (__int128)h;
        s4 += (unsigned __int128)(g * g * g * g) * h;
    }
    if (N == 0) { for (int i = 0; i < 6; i++) s[i] = 0.0; s[5] = (double)n / 4.0; return BBB_OK; }
    const double dN = (double)N;
    const double mean = (double)s1 / dN;
    const double m2 = (double)s2 / dN, m3 = (double)s3 / dN, m4 = (double)s4 / dN;
    const double var = m2 - mean * mean;
    const double c4 = m4 - 4.0 * mean * m3 + 6.0 * mean * mean * m2 - 3.0 * mean * mean * mean * mean;
    s[0] = mean;
    s[1] = var;
    s[2] = var > 0.0 ? c4 / (var * var) - 3.0 : 0.0;
    // chi-square vs N(0, n/4) integrated over unit-width bins centred on the integers
    const double sigma = std::sqrt((double)n / 4.0);
    double chi_n = 0.0, chi_b = 0.0;
    for (uint32_t b = 0; b < bins; b++) {
        double g = (double)b - half;
        double p = 0.5 * (std::erf((g + 0.5) / (sigma * M_SQRT2)) - std::erf((g - 0.5) / (sigma * M_SQRT2)));
        double ex = p * dN;
        if (ex > 0.0) { double d = (double)hist[b] - ex; chi_n += d * d / ex; }
    }
    // exact law: g + n/2 ~ Binomial(n, 1/2) (wrapped bin n -> 0 in gateware mode)
    for (uint32_t b = 0; b < bins; b++) {
        double lp = std::lgamma((double)n + 1.0) - std::lgamma((double)b + 1.0) -
                    std::lgamma((double)(n - b) + 1.0) - (double)n * M_LN2;
        double p = std::exp(lp);
        if (model == BBB_MODEL_GATEWARE && b == 0) p += std::exp(-(double)n * M_LN2);
        double ex = p * dN;
        if (ex > 0.0) { double d = (double)hist[b] - ex; chi_b += d * d / ex; }
    }
    s[3] = chi_n;
    s[4] = chi_b;
    s[5] = (double)n / 4.0;
    return BBB_OK;
}

}  // extern "C"

// ---------------------------------------------------------------------------
// LOP3 issue-rate micro-benchmark (roofline denominator for the fused kernels)
// ---------------------------------------------------------------------------
__global__ void __launch_bounds__(256) lop3_peak_kernel(uint32_t *out, uint32_t iters, uint32_t seed)
{
    uint32_t a[16];
#pragma unroll
    for (int i = 0; i < 16; i++) a[i] = seed * (threadIdx.x + 1) + i * 0x9E3779B9u;
    uint32_t k1 = seed ^ 0x5bd1e995u, k2 = seed + 0x1234567u;
    for (uint32_t it = 0; it < iters; it++) {
#pragma unroll
        for (int r = 0; r < 8; r++) {
#pragma unroll
            for (int i = 0; i < 16; i++) {
                // one LOP3 (a ^ b ^ c is a single three-input LUT) per statement, 16 independent chains
                asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(a[i]) : "r"(k1), "r"(k2));
            }
        }
    }
    uint32_t acc = 0;
#pragma unroll
    for (int i = 0; i < 16; i++) acc ^= a[i];
    if (acc == 0xdeadbeefu) out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
}

extern "C" int bbb_measure_lop3_peak(bbb_ctx *c, double *out)
{
    if (!c || !out) return BBB_EINVAL;
    BBB_CUDA(c, cudaSetDevice(c->device));
    uint32_t *d = (uint32_t *)c->scratch(7, (size_t)c->sm_count * 8 * 256 * sizeof(uint32_t));
    if (!d) return BBB_ECUDA;
    const uint32_t iters = 20000;
    const int blocks = c->sm_count * 8;
    lop3_peak_kernel<<<blocks, 256, 0, c->stream>>>(d, 200, 1);   // warm-up
    c->begin_call();
    lop3_peak_kernel<<<blocks, 256, 0, c->stream>>>(d, iters, 1);
    c->last_launches = 1;
    c->end_call();
    BBB_CUDA(c, cudaStreamSynchronize(c->stream));
    BBB_CUDA(c, cudaGetLastError());
    double ms = bbb_last_device_ms(c);
    double ops = (double)blocks * 256.0 * (double)iters * 128.0;
    *out = ops / (ms * 1e-3);
    c->last_kernel = "lop3_peak";
    return BBB_OK;
}
