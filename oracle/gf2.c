/*
 * oracle/gf2.c -- CPU ORACLE (test infrastructure, NOT the product path).
 *
 * Restatement in C of the reference's Rust GF(2) library.  Every function
 * names the Rust lines it follows (paths relative to
 * /root/reference/software/rnghunt/src/).  Rust `assert!`/index panics become
 * abort() so that a misuse is as loud here as it is there.
 */
#include "gf2.h"

#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#define GF2_ASSERT(cond)                                                        \
    do {                                                                        \
        if (!(cond)) {                                                          \
            fprintf(stderr, "oracle/gf2.c:%d: assertion failed: %s\n", __LINE__, #cond); \
            abort();                                                            \
        }                                                                       \
    } while (0)

static uint64_t *xcalloc(size_t nwords)
{
    uint64_t *p = (uint64_t *)calloc(nwords ? nwords : 1, sizeof(uint64_t));
    if (!p) abort();
    return p;
}

/* ======================================================================== */
/* BinaryVector                                                             */
/* ======================================================================== */

bvec bv_alloc(size_t n, size_t len)
{
    bvec v = { n, len, xcalloc(len) };
    return v;
}

void bv_free(bvec *v)
{
    free(v->data);
    v->data = NULL;
    v->n = v->len = 0;
}

bvec bv_clone(const bvec *v)
{
    bvec o = bv_alloc(v->n, v->len);
    memcpy(o.data, v->data, v->len * sizeof(uint64_t));
    return o;
}

/* binary_vector.rs:21-33 */
int bv_get(const bvec *v, size_t i)
{
    GF2_ASSERT(i < v->n);
    GF2_ASSERT(i / 64 < v->len);
    return (int)((v->data[i / 64] >> (63 - (i % 64))) & 1);
}

/* binary_vector.rs:221-223 */
bvec bv_from_words(size_t n, const uint64_t *w, size_t len)
{
    bvec o = bv_alloc(n, len);
    memcpy(o.data, w, len * sizeof(uint64_t));
    return o;
}

/* binary_vector.rs:229-239 */
bvec bv_from_bits(const uint8_t *bits, size_t n)
{
    bvec o = bv_alloc(n, gf2_numwords(n));
    for (size_t idx = 0; idx < n; idx++) {
        GF2_ASSERT(bits[idx] == 0 || bits[idx] == 1);
        if (bits[idx] == 1) o.data[idx / 64] |= (uint64_t)1 << (63 - (idx % 64));
    }
    return o;
}

/* binary_vector.rs:242-253 */
bvec bv_from_bitstring(const char *s)
{
    size_t n = strlen(s);
    uint8_t *bits = (uint8_t *)malloc(n ? n : 1);
    for (size_t i = 0; i < n; i++) {
        GF2_ASSERT(s[i] == '0' || s[i] == '1');
        bits[i] = (uint8_t)(s[i] - '0');
    }
    bvec o = bv_from_bits(bits, n);
    free(bits);
    return o;
}

/* binary_vector.rs:256-264 */
void bv_to_bits(const bvec *v, uint8_t *out)
{
    for (size_t i = 0; i < v->n; i++) out[i] = (uint8_t)bv_get(v, i);
}

/* binary_vector.rs:267-277 */
void bv_to_bitstring(const bvec *v, char *out)
{
    for (size_t i = 0; i < v->n; i++) out[i] = bv_get(v, i) ? '1' : '0';
    out[v->n] = '\0';
}

/* binary_vector.rs:280-304 */
bvec bv_slice(const bvec *v, size_t start, size_t end)
{
    GF2_ASSERT(end >= start);
    GF2_ASSERT(end <= v->n);
    size_t n = end - start;
    size_t nwords = gf2_numwords(n);
    size_t offset = start % 64;
    if (offset == 0) {
        /* self.data[start/64..(end+63)/64] */
        size_t lo = start / 64, hi = (end + 63) / 64;
        GF2_ASSERT(lo <= hi && hi <= v->len);
        return bv_from_words(n, v->data + lo, hi - lo);
    }
    bvec o = bv_alloc(n, nwords);
    for (size_t idx = 0; idx < nwords; idx++) {
        size_t wordidx = (start + idx * 64) / 64;
        GF2_ASSERT(wordidx < v->len);
        o.data[idx] = v->data[wordidx] << offset;
        if (idx != (nwords - 1) || wordidx != (end - 1) / 64) {
            GF2_ASSERT(wordidx + 1 < v->len);
            o.data[idx] |= v->data[wordidx + 1] >> (64 - offset);
        }
    }
    return o;
}

/* binary_vector.rs:307-314 */
size_t bv_firstbit(const bvec *v)
{
    for (size_t i = 0; i < v->n; i++)
        if (bv_get(v, i)) return i;
    return v->n;
}

/* binary_vector.rs:317-323 (counts every stored word, padding included) */
uint32_t bv_count_ones(const bvec *v)
{
    uint32_t c = 0;
    for (size_t i = 0; i < v->len; i++) c += (uint32_t)__builtin_popcountll(v->data[i]);
    return c;
}

/* binary_vector.rs:41-102: the six bitwise operators loop over numwords(self.n) */
#define GF2_BITOP(NAME, OP)                                                     \
    void NAME##_assign(bvec *a, const bvec *b)                                  \
    {                                                                           \
        GF2_ASSERT(a->n == b->n);                                               \
        size_t nw = gf2_numwords(a->n);                                         \
        GF2_ASSERT(nw <= a->len && nw <= b->len);                               \
        for (size_t w = 0; w < nw; w++) a->data[w] OP b->data[w];               \
    }                                                                           \
    bvec NAME(const bvec *a, const bvec *b)                                     \
    {                                                                           \
        bvec o = bv_clone(a);                                                   \
        NAME##_assign(&o, b);                                                   \
        return o;                                                               \
    }
GF2_BITOP(bv_and, &=)
GF2_BITOP(bv_or, |=)
GF2_BITOP(bv_xor, ^=)

/* binary_vector.rs:104-124 */
bvec bv_shl(const bvec *v, size_t i)
{
    GF2_ASSERT(i <= v->n);
    size_t n = v->n - i;
    size_t nwords = gf2_numwords(n);
    size_t offset = i % 64;
    if (offset == 0) {
        GF2_ASSERT(i / 64 <= v->len);
        return bv_from_words(n, v->data + i / 64, v->len - i / 64);
    }
    bvec o = bv_alloc(n, nwords);
    for (size_t idx = 0; idx < nwords; idx++) {
        size_t wordidx = (i + idx * 64) / 64;
        GF2_ASSERT(wordidx < v->len);
        o.data[idx] = v->data[wordidx] << offset;
        if (idx != (nwords - 1) || wordidx != (v->n - 1) / 64) {
            GF2_ASSERT(wordidx + 1 < v->len);
            o.data[idx] |= v->data[wordidx + 1] >> (64 - offset);
        }
    }
    return o;
}

/* binary_vector.rs:126-144 */
void bv_shl_assign(bvec *v, size_t i)
{
    GF2_ASSERT(i <= v->n);
    v->n -= i;
    size_t offset = i % 64;
    if (offset == 0) {
        size_t drop = i / 64;
        GF2_ASSERT(drop <= v->len);
        memmove(v->data, v->data + drop, (v->len - drop) * sizeof(uint64_t));
        v->len -= drop;
    } else {
        size_t nwords = gf2_numwords(v->n + i);
        GF2_ASSERT(nwords <= v->len);
        for (size_t idx = 0; idx < nwords; idx++) {
            v->data[idx] <<= offset;
            if (idx != nwords - 1) v->data[idx] |= v->data[idx + 1] >> (64 - offset);
        }
        v->len = nwords; /* truncate(nwords) */
    }
}

/* binary_vector.rs:146-166 */
void bv_shr_assign(bvec *v, size_t i)
{
    v->n += i;
    size_t nwords = gf2_numwords(v->n);
    size_t offset = i % 64;
    size_t lead = i / 64;
    size_t newlen = lead + v->len;
    if (nwords > newlen) newlen += 1;       /* data.push(0) */
    uint64_t *d = xcalloc(newlen);
    memcpy(d + lead, v->data, v->len * sizeof(uint64_t));
    free(v->data);
    v->data = d;
    v->len = newlen;
    if (offset != 0) {
        GF2_ASSERT(nwords <= v->len);
        for (size_t idx = nwords; idx-- > lead;) {
            v->data[idx] >>= offset;
            if (idx != 0) v->data[idx] |= v->data[idx - 1] << (64 - offset);
        }
    }
}

/* binary_vector.rs:168-192 */
bvec bv_shr(const bvec *v, size_t i)
{
    size_t n = v->n + i;
    size_t nwords = gf2_numwords(n);
    size_t offset = i % 64;
    if (offset == 0) {
        bvec o = bv_alloc(n, i / 64 + v->len);
        memcpy(o.data + i / 64, v->data, v->len * sizeof(uint64_t));
        return o;
    }
    bvec o = bv_alloc(n, nwords);
    size_t srcidx = 0;
    for (size_t dstidx = i / 64; dstidx < nwords; dstidx++, srcidx++) {
        if (srcidx > 0) {
            GF2_ASSERT(srcidx - 1 < v->len);
            o.data[dstidx] = v->data[srcidx - 1] << (64 - offset);
        }
        if (dstidx != nwords - 1 || srcidx < gf2_numwords(v->n)) {
            GF2_ASSERT(srcidx < v->len);
            o.data[dstidx] |= v->data[srcidx] >> offset;
        }
    }
    return o;
}

/* binary_vector.rs:194-213.  Restated literally, including the low-bits tail mask. */
int bv_eq(const bvec *a, const bvec *b)
{
    if (a->n != b->n) return 0;
    int eq = 1;
    size_t nwords = gf2_numwords(a->n);
    GF2_ASSERT(nwords >= 1); /* Rust: 0..(nwords-1) underflows for n == 0 */
    GF2_ASSERT(nwords <= a->len && nwords <= b->len);
    for (size_t w = 0; w + 1 < nwords; w++) eq &= a->data[w] == b->data[w];
    if (a->n % 64 != 0) {
        uint64_t mask = 0xFFFFFFFFFFFFFFFFull >> (64 - (a->n % 64));
        eq &= (a->data[nwords - 1] & mask) == (b->data[nwords - 1] & mask);
    } else {
        eq &= a->data[nwords - 1] == b->data[nwords - 1];
    }
    return eq;
}

/* ======================================================================== */
/* BinaryMatrix                                                             */
/* ======================================================================== */

/* binary_matrix.rs:27-30 */
bmat bm_from_words(size_t nrows, size_t ncols, const uint64_t *w)
{
    bmat m = { nrows, ncols, gf2_numwords(nrows), NULL };
    m.data = xcalloc(m.words_per_col * ncols);
    memcpy(m.data, w, m.words_per_col * ncols * sizeof(uint64_t));
    return m;
}

void bm_free(bmat *m)
{
    free(m->data);
    m->data = NULL;
}

/* binary_matrix.rs:33-37 */
bvec bm_col(const bmat *m, size_t col)
{
    GF2_ASSERT(col < m->ncols);
    return bv_from_words(m->nrows, m->data + col * m->words_per_col, m->words_per_col);
}

/* binary_matrix.rs:40-50 */
bvec bm_row(const bmat *m, size_t row)
{
    bvec o = bv_alloc(m->ncols, gf2_numwords(m->ncols));
    for (size_t col = 0; col < m->ncols; col++) {
        bvec c = bm_col(m, col);
        if (bv_get(&c, row)) o.data[col / 64] |= (uint64_t)1 << (63 - (col % 64));
        bv_free(&c);
    }
    return o;
}

/* binary_matrix.rs:53-64 */
bvec bm_dot(const bmat *m, const bvec *x)
{
    GF2_ASSERT(x->n == m->ncols);
    bvec o = bv_alloc(m->nrows, m->words_per_col);
    for (size_t col = 0; col < m->ncols; col++) {
        if (bv_get(x, col)) {
            for (size_t idx = 0; idx < m->words_per_col; idx++)
                o.data[idx] ^= m->data[col * m->words_per_col + idx];
        }
    }
    return o;
}

/* binary_matrix.rs:68-76 */
bvec bm_recur(const bmat *m, const bvec *x0, size_t n)
{
    uint8_t *bits = (uint8_t *)malloc(n ? n : 1);
    bvec x = bv_clone(x0);
    for (size_t t = 0; t < n; t++) {
        bvec y = bm_dot(m, &x);
        bv_free(&x);
        x = y;
        bits[t] = (uint8_t)bv_get(&x, 0);
    }
    bvec o = bv_from_bits(bits, n);
    bv_free(&x);
    free(bits);
    return o;
}

/* binary_matrix.rs:81-101.  `draw` replaces rand 0.3's Range::ind_sample / Rng::choose. */
bmat bm_random(size_t nrows, size_t ncols, const size_t *rowweights, size_t nweights,
               gf2_draw_fn draw, void *ctx)
{
    bmat m = { nrows, ncols, gf2_numwords(nrows), NULL };
    m.data = xcalloc(m.words_per_col * ncols);
    size_t *colweights = (size_t *)calloc(ncols, sizeof(size_t));
    size_t *candidates = (size_t *)malloc(ncols * sizeof(size_t));
    size_t *chosen = (size_t *)malloc((ncols ? ncols : 1) * sizeof(size_t));
    for (size_t rowidx = 0; rowidx < nrows; rowidx++) {
        size_t weight = rowweights[draw(ctx, (uint32_t)nweights)];
        size_t nchosen = 0;
        for (size_t k = 0; k < weight; k++) {
            size_t mincw = colweights[0];
            for (size_t c = 1; c < ncols; c++)
                if (colweights[c] < mincw) mincw = colweights[c];
            size_t ncand = 0;
            for (size_t c = 0; c < ncols; c++) {
                int taken = 0;
                for (size_t j = 0; j < nchosen; j++) taken |= chosen[j] == c;
                if (colweights[c] == mincw && !taken) candidates[ncand++] = c;
            }
            GF2_ASSERT(ncand > 0); /* rng.choose(&candidates).unwrap() */
            size_t colidx = candidates[draw(ctx, (uint32_t)ncand)];
            m.data[colidx * m.words_per_col + rowidx / 64] |= (uint64_t)1 << (63 - (rowidx % 64));
            colweights[colidx] += 1;
            chosen[nchosen++] = colidx;
        }
    }
    free(colweights);
    free(candidates);
    free(chosen);
    return m;
}

/* ======================================================================== */
/* BinaryPolynomial                                                         */
/* ======================================================================== */

/* binary_polynomial.rs:51-53 */
bpoly bp_from_coefficients(const uint8_t *c, size_t n)
{
    bpoly p = { bv_from_bits(c, n) };
    return p;
}

void bp_free(bpoly *p) { bv_free(&p->c); }

bpoly bp_clone(const bpoly *p)
{
    bpoly o = { bv_clone(&p->c) };
    return o;
}

/* binary_polynomial.rs:56-63 */
long bp_degree(const bpoly *p)
{
    for (size_t i = 0; i < p->c.n; i++)
        if (bv_get(&p->c, i)) return (long)(p->c.n - 1 - i);
    return -1;
}

/* binary_polynomial.rs:69-85 */
int bp_eval(const bpoly *p, const bvec *x)
{
    GF2_ASSERT((long)x->n == bp_degree(p) + 1);
    GF2_ASSERT(x->n == p->c.n);
    uint8_t parity = 0;
    for (size_t idx = 0; idx < gf2_numwords(x->n); idx++) {
        GF2_ASSERT(idx < p->c.len && idx < x->len);
        uint64_t v = p->c.data[idx] & x->data[idx];
        v ^= v >> 32;
        v ^= v >> 16;
        v ^= v >> 8;
        v ^= v >> 4;
        parity ^= (uint8_t)((0x6996ull >> (v & 0xf)) & 1);
    }
    return parity;
}

/* binary_polynomial.rs:88-133 */
bpoly bp_modmult(const bpoly *f, const bpoly *g, const bpoly *p)
{
    GF2_ASSERT(f->c.n == g->c.n);
    GF2_ASSERT(f->c.n == p->c.n);
    long dp = bp_degree(p);
    GF2_ASSERT(dp >= 0); /* `as usize` of -1 would index out of range below */
    size_t degree_p = (size_t)dp;

    bpoly gs = bp_clone(g);
    bpoly r = bp_clone(p);
    bv_xor_assign(&r.c, &p->c);

    long top = bp_degree(f) + 2;
    for (long bit = 1; bit < top; bit++) {
        if (bv_get(&f->c, f->c.n - (size_t)bit)) bv_xor_assign(&r.c, &gs.c);
        bv_shl_assign(&gs.c, 1);
        gs.c.n += 1;
        if (bv_get(&gs.c, gs.c.n - degree_p - 1)) bv_xor_assign(&gs.c, &p->c);
    }
    bp_free(&gs);
    return r;
}

static bpoly bp_monomial(size_t n, size_t pos_from_end)
{
    uint8_t *bits = (uint8_t *)calloc(n ? n : 1, 1);
    GF2_ASSERT(n >= pos_from_end);
    bits[n - pos_from_end] = 1;
    bpoly o = bp_from_coefficients(bits, n);
    free(bits);
    return o;
}

/* binary_polynomial.rs:136-163 */
bpoly bp_modexp(const bpoly *p, const bvec *k)
{
    size_t fb = bv_firstbit(k);
    if (fb == k->n) return bp_monomial(p->c.n, 1);      /* the polynomial 1 */

    bpoly x = bp_monomial(p->c.n, 2);                   /* the polynomial x */
    bpoly f = bp_clone(&x);
    for (size_t bit = fb + 1; bit < k->n; bit++) {
        bpoly t = bp_modmult(&f, &f, p);
        bp_free(&f);
        f = t;
        if (bv_get(k, bit)) {
            t = bp_modmult(&f, &x, p);
            bp_free(&f);
            f = t;
        }
    }
    bp_free(&x);
    return f;
}

/* binary_polynomial.rs:166-175: compares the data words (Vec equality) only */
int bp_check_integer(const bpoly *p, const bvec *k)
{
    bpoly f = bp_modexp(p, k);
    bpoly one = bp_monomial(p->c.n, 1);
    int same = f.c.len == one.c.len &&
               memcmp(f.c.data, one.c.data, one.c.len * sizeof(uint64_t)) == 0;
    bp_free(&f);
    bp_free(&one);
    return same;
}

/* binary_polynomial.rs:178-212 */
int bp_is_primitive(const bpoly *p)
{
    long deg = bp_degree(p);
    if (deg == -1) return 1;
    if (!bv_get(&p->c, p->c.n - 1)) return 0;
    if (bv_count_ones(&p->c) % 2 != 1) return 0;

    bvec factors[64];
    size_t nf = gf2_get_factors((size_t)deg, factors, 64);
    int result = 1;
    if (!bp_check_integer(p, &factors[0])) {
        result = 0;
    } else {
        for (size_t i = 1; i < nf; i++) {
            if (bp_check_integer(p, &factors[i])) {
                result = 0;
                break;
            }
        }
    }
    for (size_t i = 0; i < nf; i++) bv_free(&factors[i]);
    return result;
}

/* binary_polynomial.rs:22-44 */
size_t bp_display(const bpoly *p, char *out, size_t cap)
{
    size_t pos = 0;
    int first = 1;
    if (cap) out[0] = '\0';
    for (size_t i = 0; i < p->c.n; i++) {
        if (!bv_get(&p->c, i)) continue;
        size_t degree = p->c.n - 1 - i;
        char part[32];
        if (degree == 0) snprintf(part, sizeof part, "1");
        else if (degree == 1) snprintf(part, sizeof part, "x");
        else snprintf(part, sizeof part, "x^%zu", degree);
        int w = snprintf(out + pos, pos < cap ? cap - pos : 0, "%s%s", first ? "" : " + ", part);
        pos += (size_t)w;
        first = 0;
    }
    if (first) pos += (size_t)snprintf(out, cap, "0");
    return pos;
}

/* ======================================================================== */
/* berlekamp_massey.rs:5-31                                                 */
/* ======================================================================== */
bpoly gf2_berlekamp_massey(const bvec *s)
{
    size_t sw = gf2_numwords(s->n);
    bpoly c = { bv_alloc(0, sw) };
    GF2_ASSERT(sw >= 1);
    c.c.data[0] = 0x8000000000000000ull;
    bpoly b = bp_clone(&c);
    size_t l = 0;
    long m = -1;
    for (size_t n = 0; n < s->n; n++) {
        bvec x = bv_slice(s, s->n - n - 1, s->n - n + l);
        c.c.n = l + 1;
        if (bp_eval(&c, &x) == 1) {
            bpoly t = bp_clone(&c);
            c.c.n = s->n;
            b.c.n = s->n;
            size_t offset = (size_t)((long)n - m);
            bvec sh = bv_shr(&b.c, offset);
            bvec sl = bv_slice(&sh, 0, s->n);
            bv_xor_assign(&c.c, &sl);
            bv_free(&sh);
            bv_free(&sl);
            if (l <= n / 2) {
                l = n + 1 - l;
                m = (long)n;
                bp_free(&b);
                b = t;
            } else {
                bp_free(&t);
            }
        }
        bv_free(&x);
    }
    bpoly out = { bv_slice(&c.c, 0, l + 1) };
    bp_free(&c);
    bp_free(&b);
    return out;
}

/* ======================================================================== */
/* factors.rs:7-16 over the generated table                                  */
/* ======================================================================== */
#include "mersenne_factors.inc" /* BBB_MF_OFFSETS, BBB_MF_ENTRIES, BBB_MF_WORDS */

size_t gf2_factor_count(size_t n)
{
    GF2_ASSERT(n >= 1);
    GF2_ASSERT(n <= 512);
    return BBB_MF_OFFSETS[n] - BBB_MF_OFFSETS[n - 1];
}

size_t gf2_factor_words(size_t n, size_t idx, const uint64_t **words)
{
    GF2_ASSERT(idx < gf2_factor_count(n));
    size_t e = BBB_MF_OFFSETS[n - 1] + idx;
    *words = (const uint64_t *)BBB_MF_WORDS + BBB_MF_ENTRIES[e];
    return BBB_MF_ENTRIES[e + 1] - BBB_MF_ENTRIES[e];
}

size_t gf2_get_factors(size_t n, bvec *out, size_t cap)
{
    size_t cnt = gf2_factor_count(n);
    GF2_ASSERT(cnt <= cap);
    for (size_t i = 0; i < cnt; i++) {
        const uint64_t *w;
        size_t nw = gf2_factor_words(n, i, &w);
        out[i] = bv_from_words(nw * 64, w, nw);
    }
    return cnt;
}
