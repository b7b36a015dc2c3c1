/*
 * oracle/gf2.h -- CPU ORACLE (test infrastructure, NOT the product path).
 *
 * Plain-C restatement of the reference's GF(2) library
 * (software/rnghunt/src/{binary_vector,binary_matrix,binary_polynomial,
 * berlekamp_massey,factors}.rs).  Same word layouts, same algorithms, same
 * length bookkeeping (`n` and the length of `data` are tracked separately,
 * as the Rust `Vec<u64>` does), so the results are bit-for-bit those of the
 * Rust crate.  Only tests/, __graft_entry__.smoke() and bench.py's
 * cpu_baseline / --impl reference legs may link or call this.
 *
 * Parity pin: the reference's 33 inline unit tests are ported value for value
 * in tests/test_oracle_gf2.py, and software/rnghunt/matrices/{16..512} are
 * positive known-answer fixtures (tests/golden/).
 */
#ifndef BBB_ORACLE_GF2_H
#define BBB_ORACLE_GF2_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* binary_vector.rs:13-19.  bit i = data[i/64] >> (63 - i%64) & 1 (MSbit first). */
typedef struct {
    size_t n;        /* number of valid bits                                  */
    size_t len;      /* data.len(): number of u64 words actually held         */
    uint64_t *data;
} bvec;

/* binary_matrix.rs:10-20.  Column-major, word data[col*wpc + r/64], row r at bit 63-r%64. */
typedef struct {
    size_t nrows, ncols, words_per_col;
    uint64_t *data;
} bmat;

/* binary_polynomial.rs:9-11.  coefficient index 0 = highest power. */
typedef struct {
    bvec c;
} bpoly;

/* lib.rs:18-21 */
static inline size_t gf2_numwords(size_t n) { return (n + 63) / 64; }

/* ---- BinaryVector ------------------------------------------------------- */
bvec bv_alloc(size_t n, size_t len);                 /* zeroed                */
void bv_free(bvec *v);
bvec bv_clone(const bvec *v);
int  bv_get(const bvec *v, size_t i);                /* Index, :21-33         */
bvec bv_from_words(size_t n, const uint64_t *w, size_t len);   /* :221-223    */
bvec bv_from_bits(const uint8_t *bits, size_t n);    /* :229-239              */
bvec bv_from_bitstring(const char *s);               /* :242-253              */
void bv_to_bits(const bvec *v, uint8_t *out);        /* :256-264              */
void bv_to_bitstring(const bvec *v, char *out);      /* :267-277 (NUL added)  */
bvec bv_slice(const bvec *v, size_t start, size_t end);  /* :280-304          */
size_t bv_firstbit(const bvec *v);                   /* :307-314              */
uint32_t bv_count_ones(const bvec *v);               /* :317-323              */
bvec bv_and(const bvec *a, const bvec *b);           /* :41-51                */
void bv_and_assign(bvec *a, const bvec *b);          /* :53-60                */
bvec bv_or(const bvec *a, const bvec *b);            /* :62-72                */
void bv_or_assign(bvec *a, const bvec *b);           /* :74-81                */
bvec bv_xor(const bvec *a, const bvec *b);           /* :83-93                */
void bv_xor_assign(bvec *a, const bvec *b);          /* :95-102               */
bvec bv_shl(const bvec *v, size_t i);                /* :104-124              */
void bv_shl_assign(bvec *v, size_t i);               /* :126-144              */
bvec bv_shr(const bvec *v, size_t i);                /* :168-192              */
void bv_shr_assign(bvec *v, size_t i);               /* :146-166              */
int  bv_eq(const bvec *a, const bvec *b);            /* :194-213              */

/* ---- BinaryMatrix ------------------------------------------------------- */
bmat bm_from_words(size_t nrows, size_t ncols, const uint64_t *w);  /* :27-30 */
void bm_free(bmat *m);
bvec bm_col(const bmat *m, size_t col);              /* :33-37                */
bvec bm_row(const bmat *m, size_t row);              /* :40-50                */
bvec bm_dot(const bmat *m, const bvec *x);           /* :53-64                */
bvec bm_recur(const bmat *m, const bvec *x, size_t n);   /* :68-76            */
/* binary_matrix.rs:81-101 with the RNG abstracted: `draw(ctx, m)` must return a
 * uniform index in [0, m).  The reference uses an OS-seeded thread_rng (parity
 * unpinned); the build DEFINES the draw sequence in oracle/rnghunt_ref.c. */
typedef uint32_t (*gf2_draw_fn)(void *ctx, uint32_t m);
bmat bm_random(size_t nrows, size_t ncols, const size_t *rowweights, size_t nweights,
               gf2_draw_fn draw, void *ctx);

/* ---- BinaryPolynomial --------------------------------------------------- */
bpoly bp_from_coefficients(const uint8_t *c, size_t n);   /* :51-53           */
void  bp_free(bpoly *p);
bpoly bp_clone(const bpoly *p);
long  bp_degree(const bpoly *p);                     /* :56-63                */
int   bp_eval(const bpoly *p, const bvec *x);        /* :69-85                */
bpoly bp_modmult(const bpoly *f, const bpoly *g, const bpoly *p);  /* :88-133 */
bpoly bp_modexp(const bpoly *p, const bvec *k);      /* :136-163              */
int   bp_check_integer(const bpoly *p, const bvec *k);   /* :166-175          */
int   bp_is_primitive(const bpoly *p);               /* :178-212              */
/* fmt::Display, :22-44.  Returns number of chars written (excluding NUL). */
size_t bp_display(const bpoly *p, char *out, size_t cap);

/* ---- berlekamp_massey.rs:5-31 ------------------------------------------- */
bpoly gf2_berlekamp_massey(const bvec *s);

/* ---- factors.rs:7-16 ---------------------------------------------------- */
/* Returns the number of entries for 2^n-1 ([r, r/q1, ...]); entry i is written
 * as a fresh bvec (n = 64*words, MS word first) into out[i] (cap entries). */
size_t gf2_get_factors(size_t n, bvec *out, size_t cap);
/* Raw table access (also used to ship the exponents to the GPU). */
size_t gf2_factor_count(size_t n);
size_t gf2_factor_words(size_t n, size_t idx, const uint64_t **words);

#ifdef __cplusplus
}
#endif
#endif
