/*
 * bbb_b200.h -- C ABI of the B200-native noise / test-pattern datapath of
 * adamgreig/basebandboard (libbbb_b200.so, hand-written CUDA for sm_100a).
 *
 * The reference has no FFI of its own: its hot path is the Rust
 * crate software/rnghunt (lib.rs:12-21), the gateware generator blocks
 * gateware/bbb/rng.py / prbs.py (whose numpy/Python models are the software
 * contract) and software/clt-grng/clt-grng-evaluate.py.  Each entry point below
 * names the reference interface it stands in for.  A Rust host binds these
 * with a plain `extern "C"` block; the Python host mirror
 * in basebandboard_b200/ binds them with ctypes.
 *
 * Conventions
 *   - every call returns 0 on success or a negative BBB_E* code;
 *     bbb_last_error(ctx) returns the message of the last failure on that ctx.
 *   - one ctx per GPU per host thread; calls on one ctx are serialised by the
 *     caller; no global mutable state; nothing throws across the boundary.
 *   - generator state / URNG words: little-endian array of u64, state bit i is
 *     word i/64 bit i%64 (bit i has weight 2^i, rng.py:135).
 *   - recurrence matrix: CSR, row r is the XOR of x[taps[row_ptr[r]..row_ptr[r+1])]
 *     (LUTOPT.from_packed, rng.py:42-55).
 *   - rnghunt structures use the crate's layout: BinaryMatrix column-major,
 *     word data[col*wpc + r/64], row r at bit 63 - r%64 (binary_matrix.rs:15-19);
 *     polynomials MSbit-first with coefficient 0 = highest power
 *     (binary_polynomial.rs:9-11, binary_vector.rs:25-31).
 *   - sample indexing (DESIGN.md "definitions"): g[i] = clt(A^(i+1) . x0), i >= 0;
 *     URNG word block i = A^(i+1) . x0; PRBS bit i = output of the i-th clock
 *     from state0 (prbs.py:112-119).
 *   - `*_dev` pointers are device pointers on the ctx's GPU (e.g. a torch
 *     tensor's data_ptr); all other pointers are host memory.  The host-buffer
 *     stream calls (bbb_prbs_stream, bbb_clt_stream[_dec], bbb_lutopt_stream,
 *     bbb_tx_stream) generate in chunks and overlap generation, the device-to-
 *     host copy and the host-side copy; page-locked caller memory is detected
 *     and written directly (about 56 GB/s instead of 21-42 GB/s).
 *   - there is NO CPU fallback: with no usable GPU bbb_create fails.
 */
#ifndef BBB_B200_H
#define BBB_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define BBB_OK            0
#define BBB_EINVAL       -1   /* bad argument                                   */
#define BBB_ECUDA        -2   /* CUDA runtime error (message has the detail)    */
#define BBB_ENOMEM       -3
#define BBB_ESTATE       -4   /* call order (e.g. no recurrence loaded)         */
#define BBB_EUNSUPPORTED -5

#define BBB_MODEL_GATEWARE 0  /* CLTGRNG output truncated to log2(n) signed bits (rng.py:78,108) */
#define BBB_MODEL_NUMPY    1  /* unwrapped value of the numpy model (rng.py:173-181)             */

/* how far a candidate got through rnghunt.rs:36-44 / binary_polynomial.rs:178-212 */
#define BBB_STAGE_DEGREE    0 /* Berlekamp-Massey degree != n (rnghunt.rs:39-41)          */
#define BBB_STAGE_FILTER    1 /* zero constant term or even weight (:185-192)             */
#define BBB_STAGE_ORDER     2 /* x^(2^n-1) != 1 (:198)                                    */
#define BBB_STAGE_FACTOR    3 /* x^((2^n-1)/q) == 1 for a prime q (:204-208)              */
#define BBB_STAGE_PRIMITIVE 4 /* accepted: "Found suitable recurrence matrix" (rnghunt.rs:44-45) */
#define BBB_NSTAGES         5

typedef struct bbb_ctx bbb_ctx;

/* ---- context ----------------------------------------------------------- */
int  bbb_version(void);
/* Number of CUDA devices visible to the process (0 if none / no driver): what a host sizes its pool of
 * worker threads with, one ctx each -- the analogue of num_cpus::get() in rnghunt.rs:17. */
int  bbb_device_count(void);
/* Binds `device` (cudaSetDevice ordinal), creates the ctx's stream. */
int  bbb_create(int device, bbb_ctx **out);
void bbb_destroy(bbb_ctx *ctx);
const char *bbb_last_error(const bbb_ctx *ctx);
/* Device time (ms, CUDA events on the ctx's stream) of the kernels launched by
 * the most recent call on this ctx, and how many of this library's kernels
 * that call launched. */
double   bbb_last_device_ms(const bbb_ctx *ctx);
uint32_t bbb_last_launches(const bbb_ctx *ctx);
/* Name of the kernel family the last CLT/LUTOPT call dispatched to
 * ("clt256_fused", "generic", ...). */
const char *bbb_last_kernel(const bbb_ctx *ctx);
/* Block until everything queued on the ctx's stream has finished. */
int  bbb_sync(bbb_ctx *ctx);
/* Host<->device bytes the most recent call on this ctx copied (jump-ahead
 * operators, seeds, results): what bench.py reports as h2d/d2h bytes per step. */
uint64_t bbb_last_h2d_bytes(const bbb_ctx *ctx);
uint64_t bbb_last_d2h_bytes(const bbb_ctx *ctx);
/* Kernel rounds the last bbb_prbs_detect needed (1 = every chunk's start state was right). */
uint32_t bbb_last_rounds(const bbb_ctx *ctx);
/* Tuning knobs, for tests and experiments (results never depend on them):
 *   "clt_no_spec" (0/1), "clt_spec_variant" (1: the round-1 two-pass n256 kernels, 2: one-pass), "clt_spec_min_steps",
 *   "clt_spec_max_blocks", "dec_full_stream" (0/1: decimated samples by full stream + pick), "slicer_chunk_bits",
 *   "prbs_chunk_words", "prbs_stream_reps", "hunt_batch", "hunt_slow_check" (0/1),
 *   "hunt_smem_taps" (0 auto = warp per candidate, K clocks per pass; 1 thread-per-candidate kernel with
 *   shared-memory taps for every n; 2 warp per candidate, one clock per pass), "hunt_pow_k" (K: 8 or 4),
 *   "hunt_cpw" (candidates per warp pass of the one-clock warp kernel: 32 or 16),
 *   "stream_chunk_bytes", "host_copy_threads" (host-buffer stream calls),
 *   "words_variant" (raw-word stream kernel: 0 default clt256_words7; 2 clt256_words2; 4 ... 16 the
 *   measured alternatives of DESIGN.md 4.1), "words_b192" (0/1), "seed_single" (0/1: never the four-groups-per-warp seeding kernel). */
int  bbb_set_tuning(bbb_ctx *ctx, const char *key, int64_t value);
/* Region timer: CUDA events on the ctx's stream around any number of calls.
 * bbb_timer_stop waits for the stream and returns the device time in ms. */
int  bbb_timer_start(bbb_ctx *ctx);
int  bbb_timer_stop(bbb_ctx *ctx, double *ms);

/* ---- LUTOPT uniform RNG (gateware/bbb/rng.py:14-55) --------------------- */
/* LUTOPT.from_packed(packed) (rng.py:42-55): n in [8,512], multiple of 8. */
int bbb_lutopt_load(bbb_ctx *ctx, uint32_t n, const uint32_t *row_ptr, const uint16_t *taps);
/* rng_recurrences.nN for N in {16,32,64,128,192,256} plus matrices/512. */
int bbb_lutopt_load_shipped(bbb_ctx *ctx, uint32_t n);
/* Copies the loaded recurrence out (sizes: n+1 and nnz); either may be NULL. */
int bbb_lutopt_get(const bbb_ctx *ctx, uint32_t *n, uint32_t *nnz, uint32_t *row_ptr, uint16_t *taps);
/* x_out = A^e . x_in (host-side GF(2) jump-ahead; e == 1 is one clock, rng.py:38-40). */
int bbb_lutopt_jump(bbb_ctx *ctx, const uint64_t *x_in, uint64_t e, uint64_t *x_out);
/* URNG words of clocks first+1 .. first+count (the `x` output of LUTOPT each
 * clock, rng.py:27): out[i*nw + w], nw = ceil(n/64). */
int bbb_lutopt_stream(bbb_ctx *ctx, const uint64_t *x0, uint64_t first, uint64_t count,
                      uint64_t *out_words);
int bbb_lutopt_stream_dev(bbb_ctx *ctx, const uint64_t *x0, uint64_t first, uint64_t count,
                          uint64_t *out_words_dev);

/* ---- CLT Gaussian RNG (rng.py:58-108; clt-grng-evaluate.py:8-21) ---------- */
/* Fused generate + histogram + moments over g[first, first+count); nothing is
 * stored.  hist has n bins (gateware) or n+1 (numpy), bin = g + n/2
 * (clt-grng-evaluate.py:34: bins=n, range=(-n/2, n/2)); counts are ADDED to
 * hist[], *sum (sum g) and *sumsq (sum g^2), any of which may be NULL.
 * n must be a power of two (rng.py:73-74). */
int bbb_clt_hist(bbb_ctx *ctx, const uint64_t *x0, uint64_t first, uint64_t count, uint32_t model,
                 uint64_t *hist, int64_t *sum, uint64_t *sumsq);
/* Same, result left on the device for an NCCL reduction: acc_dev[0..n] = hist
 * (n+1 slots, slot n unused in gateware mode), acc_dev[n+1] = sum (two's
 * complement), acc_dev[n+2] = sumsq.  ADDS into acc_dev; asynchronous on the
 * ctx's stream (bbb_sync to wait). */
int bbb_clt_hist_dev(bbb_ctx *ctx, const uint64_t *x0, uint64_t first, uint64_t count, uint32_t model,
                     uint64_t *acc_dev);
/* bbb_clt_hist_dev without any allocation.  No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (CUDA graph capture around the call, tight
 * serving loops, harnesses that forbid it).  Same arguments, same result (ADDS into acc_dev, asynchronous
 * on the ctx's stream), computed by the data-driven kernel for any loaded recurrence; it calls no
 * cudaMalloc / cudaFree: its device scratch is one file-scope static array of BBB_STATIC_SCRATCH_BYTES
 * that every GPU loading the library holds until exit.  Limits: count <= 2^BBB_STATIC_MAX_LOG2_COUNT
 * samples per call (BBB_EINVAL beyond that, with a message naming the limit), n <= 512.  All calls on a
 * device share the array: it keeps the last call's contents (the call clears what must start at zero),
 * and calls that overlap on the GPU -- two contexts, or two host threads, on one device -- race. */
#define BBB_STATIC_SCRATCH_BYTES 4726784u
#define BBB_STATIC_MAX_LOG2_COUNT 32u
int bbb_clt_hist_dev_static(bbb_ctx *ctx, const uint64_t *x0, uint64_t first, uint64_t count, uint32_t model,
                            uint64_t *acc_dev);
/* Materialised samples: out8[i] = g[first+i] as int8 (n <= 256), or
 * out16[i] = g[first+i] * noise_var as int16 (tx.py:75-76; noise_var = 1 gives
 * raw g).  Exactly one of out8/out16 is non-NULL. */
int bbb_clt_stream(bbb_ctx *ctx, const uint64_t *x0, uint64_t first, uint64_t count, uint32_t model,
                   uint32_t noise_var, int8_t *out8, int16_t *out16);
int bbb_clt_stream_dev(bbb_ctx *ctx, const uint64_t *x0, uint64_t first, uint64_t count, uint32_t model,
                       uint32_t noise_var, int8_t *out8_dev, int16_t *out16_dev);
/* Decimated samples: out8[j] = g[first + j*spb + delay], j < count -- the one noise sample
 * of each bit period the receiver keeps (gateware/bbb/rx.py:18-19: sample `delay` of the
 * spb samples of a bit; noise from tx.py:69-76).  Gateware widths.  The generator is
 * clocked through every sample; the CLT tree runs on the kept ones only. */
int bbb_clt_stream_dec(bbb_ctx *ctx, const uint64_t *x0, uint64_t first, uint64_t count, uint32_t spb,
                       uint32_t delay, int8_t *out8);
int bbb_clt_stream_dec_dev(bbb_ctx *ctx, const uint64_t *x0, uint64_t first, uint64_t count, uint32_t spb,
                           uint32_t delay, int8_t *out8_dev);
/* CLTGRNG of ONE URNG word x (rng.py:96-108), pure host arithmetic.  For the hardware's start-up: after reset
 * CLTGRNG.x shows log2(n) zeros (reset registers), then clt(x0), then g[0], g[1], ... (rng.py:70-72,189-190). */
int bbb_clt_value(uint32_t n, const uint64_t *x, uint32_t model, int32_t *g);
/* clt-grng-evaluate.py:18-21 from an integer histogram, in f64 with a fixed
 * summation order: stats[0]=mean, [1]=variance (ddof=0), [2]=excess kurtosis,
 * [3]=chi-square against N(0, n/4) bin probabilities, [4]=chi-square against the
 * exact Binomial(n,1/2) law, [5]=theoretical variance n/4.  Pure host arithmetic. */
int bbb_clt_stats(uint32_t n, uint32_t model, const uint64_t *hist, double *stats6);

/* ---- PRBS (gateware/bbb/prbs.py) ---------------------------------------- */
/* prbs.py:14 TAPS; returns 0 for an unsupported k. */
uint32_t bbb_prbs_tap(uint32_t k);
/* Packed PRBS-k bits [first_bit, first_bit+nbits) of the generator started at
 * state0 (prbs.py:23-35 / :112-113): bit i -> out[(i-first_bit)/64] bit (i-first_bit)%64.
 * nbits must be a multiple of 64. */
int bbb_prbs_stream(bbb_ctx *ctx, uint32_t k, uint32_t state0, uint64_t first_bit, uint64_t nbits,
                    uint64_t *out_words);
int bbb_prbs_stream_dev(bbb_ctx *ctx, uint32_t k, uint32_t state0, uint64_t first_bit, uint64_t nbits,
                        uint64_t *out_words_dev);

/* PRBSErrorDetector registers (prbs.py:61-99). */
typedef struct {
    uint32_t bit_in, prbs, err_sr, reload_ctr;
} bbb_det_state;

typedef struct {
    uint64_t err;           /* sum of `err` sampled after each clock edge (prbs.py:79)       */
    uint64_t err_valid;     /* sum of err & !reload (the quantity prbs.py:158-163 compares)  */
    uint64_t reload_events; /* edges at which err_count > k//2 fired (prbs.py:92-95)         */
    uint64_t reload_cycles; /* sum of `reload` sampled after each clock edge (prbs.py:99)    */
    uint64_t injected;      /* channel errors injected                                       */
} bbb_det_counters;

/* Generator(state0=1 advanced to first_bit) -> channel -> PRBSErrorDetector over
 * bits [first_bit, first_bit+nbits).  Channel: err_m == 0 none, else Bernoulli
 * 2^-err_m keyed (seed, absolute bit index).  `det` is the detector
 * state on entry (NULL = reset state, prbs.py:62,80,90) and is advanced; counters
 * are ADDED to *ctr. */
int bbb_prbs_detect(bbb_ctx *ctx, uint32_t k, uint64_t first_bit, uint64_t nbits, uint32_t err_m,
                    uint64_t seed, bbb_det_state *det, bbb_det_counters *ctr);
/* Same generator and detector, channel errors supplied by the caller on the device:
 * bit i of err_words_dev[j] set = PRBS bit first_bit + 64*j + i arrives inverted.  Replaces
 * wiring an arbitrary error pattern into PRBSErrorDetector.bit as the reference's own test
 * does (prbs.py:124-165).  ceil(nbits/64) words are read. */
int bbb_prbs_detect_errs_dev(bbb_ctx *ctx, uint32_t k, uint64_t first_bit, uint64_t nbits,
                             const uint64_t *err_words_dev, bbb_det_state *det, bbb_det_counters *ctr);
/* The board's noise loop without the pulse shaper: TX sample of bit b =
 * (b ? +amplitude : -amplitude) + g * noise_var registered as a 12-bit signed word (tx.py:75-81),
 * g = CLT-GRNG of the loaded recurrence from init x0 (tx.py:70-71), `spb` samples per bit; RX
 * slices at zero (rx.py:28: sliced = ~sample[-1]), keeps sample `delay` of every bit
 * (rx.py:31-32) and clocks PRBSErrorDetector with it (rx.py:43-46).  Bit i (absolute) uses noise
 * sample g[i*spb + delay].  nbits must be a multiple of 64; `injected` counts the slicer errors. */
int bbb_prbs_slicer_detect(bbb_ctx *ctx, uint32_t k, uint64_t first_bit, uint64_t nbits, const uint64_t *x0,
                           uint32_t amplitude, uint32_t noise_var, uint32_t spb, uint32_t delay,
                           bbb_det_state *det, bbb_det_counters *ctr);
/* ---- pulse shaper and the TX word (gateware/bbb/bitshaper.py, tx.py:52-81) ---- */
/* PRBSShaper (bitshaper.py:12-91): coeff[64] is one pulse (each value in (-256, 255], 8 per bit
 * duration; PRBSShaper.from_rcf builds the raised-cosine sets, basebandboard_b200/shaper.py restates
 * it).  As a sequence, the relation the reference test pins (bitshaper.py:150-163):
 * y[8m+4] = 2*b[m]-1 (else 0), S[t] = sum_k coeff[k]*y[t-k], b = PRBS-k from state 1, b[m<0] = 0 (the
 * shift register's reset value); the HDL's 13-clock pipeline latency is dropped.
 * out[t - 8*first_bit] = the 12-bit signed TX word (tx.py:80-81) x[t] = bit_en*S[t] + g[t]*noise_var,
 * wrapped to 12 bits, stored as int16, for t in [8*first_bit, 8*(first_bit+nbits)); g = CLT-GRNG of
 * the loaded recurrence from init x0 (x0 NULL = noise off).
 * k = BBB_TX_SRC_PULSE selects the other source of tx.py:59-65 (src_sel = 1): the Pulser (tx.py:20-31), whose
 * bit m is 1 when m % 256 == 0, through the same shaper. */
#define BBB_TX_SRC_PULSE 0
int bbb_tx_stream(bbb_ctx *ctx, uint32_t k, uint64_t first_bit, uint64_t nbits, const int16_t *coeff,
                  uint32_t bit_en, const uint64_t *x0, uint32_t noise_var, int16_t *out);
int bbb_tx_stream_dev(bbb_ctx *ctx, uint32_t k, uint64_t first_bit, uint64_t nbits, const int16_t *coeff,
                      uint32_t bit_en, const uint64_t *x0, uint32_t noise_var, int16_t *out_dev);
/* bbb_prbs_slicer_detect through the shaped channel (inter-symbol interference included): bit J
 * is decided on the sliced TX word x[8*(J+4) + delay] (its pulse peaks at tap 32), delay 0..7
 * (rx.py:18-19), and the decisions clock PRBSErrorDetector.  `injected` = wrong decisions. */
int bbb_prbs_shaped_slicer_detect(bbb_ctx *ctx, uint32_t k, uint64_t first_bit, uint64_t nbits,
                                  const int16_t *coeff, const uint64_t *x0, uint32_t noise_var, uint32_t delay,
                                  bbb_det_state *det, bbb_det_counters *ctr);
/* The state a locked detector with no reload pending has just before bit first_bit (>= 64) of that
 * channel: what a shard that does not start at bit 0 passes as `det` before it has its left
 * neighbour's end state; the caller compares the two and re-runs the shard on a mismatch
 * (basebandboard_b200/sharding.py).  Pure host arithmetic: ctx may be NULL and no GPU is needed. */
int bbb_prbs_spec_state(bbb_ctx *ctx, uint32_t k, uint64_t first_bit, uint32_t err_m, uint64_t seed,
                        bbb_det_state *det);

/* ---- rnghunt (software/rnghunt/src) ------------------------------------- */
/* The body of the worker loop, rnghunt.rs:27-44, on `ncand` caller-supplied
 * n x n matrices in BinaryMatrix layout (ncols*wpc u64 each, back to back; a
 * Rust host passes a.data.as_ptr()).  stage[c] = BBB_STAGE_*.  poly_words (may
 * be NULL) receives the Berlekamp-Massey polynomial of each candidate as
 * ceil((n+1)/64) MSbit-first words, coefficient 0 = highest power, valid when
 * the degree is n (zero otherwise). */
int bbb_rnghunt_eval(bbb_ctx *ctx, uint32_t n, uint64_t ncand, const uint64_t *matrices,
                     uint8_t *stage, uint64_t *poly_words);
/* Candidate c of `seed` -- BinaryMatrix::random(n, n, [3,4,4,4,4,4,4,4])
 * (binary_matrix.rs:81-101, rnghunt.rs:25) with the draws taken from the
 * counter-based generator DESIGN.md defines -- generated on the device and
 * returned in BinaryMatrix layout. */
int bbb_rnghunt_candidate(bbb_ctx *ctx, uint32_t n, uint64_t seed, uint64_t c, uint64_t *matrix_words);
/* Sweep candidates first, first+stride, ... (count of them) fully on the
 * device: generate -> recur -> Berlekamp-Massey -> primitivity, with stream
 * compaction between the stages.  best[] receives the (up to k) smallest
 * accepted candidate indices in ascending order, *nfound how many were written,
 * stage_counts[BBB_NSTAGES] is ADDED to; stage_out (may be NULL) gets one byte
 * per candidate. */
int bbb_rnghunt_sweep(bbb_ctx *ctx, uint32_t n, uint64_t seed, uint64_t first, uint64_t stride,
                      uint64_t count, uint32_t k, uint64_t *best, uint32_t *nfound,
                      uint64_t *stage_counts, uint8_t *stage_out);

/* The same sweep, also returning what the reference's quality pass looks at (util/verify.py:19-28: the
 * characteristic polynomial of an accepted matrix): poly_words[i*pw .. ) = the Berlekamp-Massey polynomial
 * of best[i] as pw = ceil((n+1)/64) MSbit-first words, coefficient 0 = highest power (binary_polynomial.rs:9-11),
 * weights[i] = its number of terms (103 for the shipped n = 192 matrix).  Either may be NULL, not both. */
int bbb_rnghunt_sweep_polys(bbb_ctx *ctx, uint32_t n, uint64_t seed, uint64_t first, uint64_t stride,
                            uint64_t count, uint32_t k, uint64_t *best, uint32_t *nfound,
                            uint64_t *stage_counts, uint64_t *poly_words, uint32_t *weights);

/* ---- measurement helpers ------------------------------------------------ */
/* Dependent-free LOP3 issue micro-benchmark: returns thread-instructions/s
 * (the integer-pipe roofline denominator). */
int bbb_measure_lop3_peak(bbb_ctx *ctx, double *thread_instr_per_s);

#ifdef __cplusplus
}
#endif
#endif /* BBB_B200_H */
