/*
 * oracle/gen.h -- CPU ORACLE (test infrastructure, NOT the product path).
 *
 * Restatement of the reference's generator datapath models:
 *   LUTOPT      gateware/bbb/rng.py:21-40  (numpy model :134-135)
 *   CLTGRNG     gateware/bbb/rng.py:70-108 (numpy model :173-181,
 *               software/clt-grng/clt-grng-evaluate.py:10-15)
 *   TX noise    gateware/bbb/tx.py:75-76
 *   PRBS        gateware/bbb/prbs.py:14,23-35 (model :112-113)
 *   PRBSErrorDetector  gateware/bbb/prbs.py:43-99 (protocol :124-165)
 * plus the definitions the reference leaves open and this build fixes
 * (sample indexing, error source, counters) -- see DESIGN.md "NEW definitions".
 *
 * Conventions at this interface (identical to include/bbb_b200.h):
 *   generator state  = little-endian array of u64, state bit i = word i/64 bit i%64
 *                      (bit i has weight 2^i as in rng.py:135)
 *   recurrence       = CSR: row r XORs x[taps[row_ptr[r] .. row_ptr[r+1])]
 *   sample index     g[i] = clt(A^(i+1) x0), i >= 0
 */
#ifndef BBB_ORACLE_GEN_H
#define BBB_ORACLE_GEN_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define ORC_MODEL_GATEWARE 0  /* output truncated to log2(n) signed bits (rng.py:78,108) */
#define ORC_MODEL_NUMPY    1  /* unwrapped int16 value of the numpy model (rng.py:173-181) */

/* x_out = A.x_in over GF(2); rng.py:38-40.  nwords = n/64 rounded up. */
void orc_lutopt_step(uint32_t n, const uint32_t *row_ptr, const uint16_t *taps,
                     const uint64_t *x_in, uint64_t *x_out);

/* x_out = A^e . x_in by dense square-and-multiply (independent check of jump-ahead). */
void orc_lutopt_jump(uint32_t n, const uint32_t *row_ptr, const uint16_t *taps,
                     uint64_t e, const uint64_t *x_in, uint64_t *x_out);

/* The pairwise-difference tree on one state; rng.py:96-105 / :173-181. */
int32_t orc_clt(uint32_t n, const uint64_t *x, int model);

/* Sequential run: samples g[first .. first+count).  `state` on entry must be
 * A^first . x0 (i.e. x0 itself when first == 0) and is advanced by `count`.
 * Any of hist / out8 / out16 may be NULL.  hist has n bins (gateware) or n+1
 * (numpy), bin = g + n/2.  out16 = g * noise_var (tx.py:75-76); pass 1 for raw g.
 * sum / sumsq accumulate Σg and Σg² exactly. */
void orc_clt_run(uint32_t n, const uint32_t *row_ptr, const uint16_t *taps,
                 uint64_t *state, uint64_t count, int model, uint32_t noise_var,
                 uint64_t *hist, int64_t *sum, uint64_t *sumsq,
                 int8_t *out8, int16_t *out16, uint64_t *out_words);

/* Same over [first, first+count) from x0 on `threads` host threads: each thread
 * jumps to its sub-range with orc_lutopt_jump and steps sequentially. */
void orc_clt_run_mt(uint32_t n, const uint32_t *row_ptr, const uint16_t *taps,
                    const uint64_t *x0, uint64_t first, uint64_t count, int model,
                    uint64_t *hist, int64_t *sum, uint64_t *sumsq, int threads);

/* PRBS: returns tap for k or 0 if k is not one of prbs.py:14's keys. */
uint32_t orc_prbs_tap(uint32_t k);
/* Writes nbits bits starting from `*state` (advanced); bit i -> out[i/64] bit i%64. */
void orc_prbs_bits(uint32_t k, uint32_t *state, uint64_t nbits, uint64_t *out);
/* State of the generator after `steps` clocks from `state0` (sequential). */
uint32_t orc_prbs_advance(uint32_t k, uint32_t state0, uint64_t steps);

/* NEW definition: Bernoulli(2^-m) error word for bits [64W, 64W+64) keyed by seed. */
uint64_t orc_err_word(uint64_t seed, uint64_t W, uint32_t m);

/* PRBSErrorDetector state, prbs.py:61-99 */
typedef struct {
    uint32_t bit_in, prbs, err_sr, reload_ctr;
} orc_det_state;

typedef struct {
    uint64_t err;          /* Σ err sampled after each clock edge              */
    uint64_t err_valid;    /* Σ err & !reload                                  */
    uint64_t reload_events;/* edges at which err_count > k//2 fired            */
    uint64_t reload_cycles;/* Σ reload sampled after each clock edge           */
    uint64_t injected;     /* number of injected channel errors                */
} orc_det_counters;

void orc_det_reset(uint32_t k, orc_det_state *s);
/* One clock edge with input `bit`; returns err | reload<<1 | fired<<2 sampled after the edge. */
uint32_t orc_det_clock(uint32_t k, orc_det_state *s, uint32_t bit);

/* Generator (state0) -> channel (err model: m == 0 none, else Bernoulli 2^-m keyed
 * (seed, absolute bit index first_bit+i)) -> detector, nbits cycles.  Generator and
 * detector states are advanced in place; counters are ADDED to *ctr. */
void orc_prbs_detect(uint32_t k, uint32_t *gen_state, orc_det_state *det,
                     uint64_t first_bit, uint64_t nbits, uint32_t m, uint64_t seed,
                     orc_det_counters *ctr);

/* The detector alone, clocked with nbits received bits (bit i = rx[i/64] >> i%64); `tx` (may be
 * NULL) are the transmitted bits, used only to count `injected` = popcount(rx ^ tx).  This is how
 * the reference's own test drives it (prbs.py:124-165: an arbitrary pattern wired into `bit`). */
void orc_det_run_bits(uint32_t k, orc_det_state *det, const uint64_t *rx, const uint64_t *tx,
                      uint64_t nbits, orc_det_counters *ctr);

#ifdef __cplusplus
}
#endif
#endif
