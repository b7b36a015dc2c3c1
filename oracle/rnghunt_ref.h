/*
 * oracle/rnghunt_ref.h -- CPU ORACLE (test infrastructure, NOT the product path).
 * rnghunt worker loop (software/rnghunt/src/bin/rnghunt.rs:23-57) + the
 * reproducible candidate stream this build defines.
 */
#ifndef BBB_ORACLE_RNGHUNT_REF_H
#define BBB_ORACLE_RNGHUNT_REF_H

#include "gf2.h"

#ifdef __cplusplus
extern "C" {
#endif

/* How far a candidate got through rnghunt.rs:36-44 / binary_polynomial.rs:178-212 */
#define ORC_STAGE_DEGREE    0  /* BM degree != n                                  */
#define ORC_STAGE_FILTER    1  /* zero constant term or even number of terms      */
#define ORC_STAGE_ORDER     2  /* x^(2^n-1) != 1                                  */
#define ORC_STAGE_FACTOR    3  /* x^((2^n-1)/q) == 1 for some prime q             */
#define ORC_STAGE_PRIMITIVE 4  /* accepted                                        */
#define ORC_NSTAGES         5

/* Candidate c of `seed`: binary_matrix.rs:81-101 with rowweights [3,4,4,4,4,4,4,4]. */
bmat orc_rnghunt_candidate(uint32_t n, uint64_t seed, uint64_t c);

/* Evaluate one matrix; optionally returns the BM polynomial's data words. */
int orc_rnghunt_eval(const bmat *a, uint64_t *poly_words, uint32_t poly_cap_words);

/* Evaluate candidates [first, first+count) on `threads` host threads. */
void orc_rnghunt_sweep(uint32_t n, uint64_t seed, uint64_t first, uint64_t count,
                       uint8_t *stage_out, uint64_t *stage_counts, int threads);

/* The same verdicts for an arbitrary list of candidate indices of `seed`. */
void orc_rnghunt_eval_list(uint32_t n, uint64_t seed, const uint64_t *cands, uint64_t count,
                           uint8_t *stage_out, uint64_t *stage_counts, int threads);

#ifdef __cplusplus
}
#endif
#endif
