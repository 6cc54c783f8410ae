/*
 * CPU oracle (plain C, unsigned __int128) for the BN254 pairing path of substrate-bn.
 * TEST INFRASTRUCTURE ONLY: only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * --impl reference legs may load this library.  The product (paritytech-bn_b200/) never does.
 *
 * What it is: a restatement of the reference crate's algorithm AS WRITTEN -- two 128-bit limbs
 * with a radix-2^128 Montgomery step, the 4-multiply Fq2 product with an explicit Montgomery
 * multiply by -1, the full-Fq2-multiply by xi, the multiply-by-one in cyclotomic_pow, the double
 * to_affine in pairing_batch -- so that (a) its outputs are the byte-exact parity oracle at 2^16
 * scale and (b) its timings describe the reference's CPU path (the Rust crate itself cannot be
 * built here: no cargo/rustc in the image).
 *
 * Parity status: PINNED.  tests/test_oracle_c.py checks it against the reference's golden vectors
 * (tests/golden/reference_vectors.json) and against the Python oracle on random inputs.
 *
 * All field elements are in Montgomery form (R = 2^256), fully reduced, and are stored exactly as
 * the crate's #[repr(C)] types are: little-endian 256-bit integers (U256([lo128, hi128])).
 * Constants are derived at init from q, r, xi and the BN parameter u (no literals copied);
 * the tests compare them with the reference's inline literals.
 *
 * Each function cites the reference file:line (relative to /root/reference) that it follows.
 */
#include <stddef.h>
#include <stdint.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

typedef unsigned __int128 u128;
typedef struct { u128 l[2]; } u256;          /* src/arith.rs:9-11  U256([u128; 2]) */

#define LO64(x) ((u128)(uint64_t)(x))
#define HI64(x) ((x) >> 64)

/* ------------------------------------------------------------------------------------------ */
/* 256-bit helpers: src/arith.rs:172-185 (Ord), :427-469 (adc / add_nocarry / sub_noborrow)      */
/* ------------------------------------------------------------------------------------------ */
static inline int u256_cmp(const u256 *a, const u256 *b) {
    if (a->l[1] != b->l[1]) return a->l[1] < b->l[1] ? -1 : 1;
    if (a->l[0] != b->l[0]) return a->l[0] < b->l[0] ? -1 : 1;
    return 0;
}
static inline int u256_is_zero(const u256 *a) { return (a->l[0] | a->l[1]) == 0; }
static inline int u256_eq(const u256 *a, const u256 *b) { return a->l[0] == b->l[0] && a->l[1] == b->l[1]; }
static inline void add_nocarry(u256 *a, const u256 *b) {
    u128 lo = a->l[0] + b->l[0];
    u128 c = lo < a->l[0];
    a->l[0] = lo;
    a->l[1] = a->l[1] + b->l[1] + c;
}
static inline void sub_noborrow(u256 *a, const u256 *b) {
    u128 br = a->l[0] < b->l[0];
    a->l[0] -= b->l[0];
    a->l[1] = a->l[1] - b->l[1] - br;
}
static inline void div2(u256 *a) {              /* src/arith.rs:399-405 */
    a->l[0] = (a->l[0] >> 1) | (a->l[1] << 127);
    a->l[1] >>= 1;
}

/* src/arith.rs:475-490: a + b*c + carry -> (low 128 returned, high 128 into *carry), done in
 * 64-bit halves exactly like the reference's split_u128/combine_u128 formulation. */
static inline u128 mac_with_carry(u128 a, u128 b, u128 c, u128 *carry) {
    u128 b_hi = HI64(b), b_lo = LO64(b);
    u128 c_hi = HI64(c), c_lo = LO64(c);
    u128 a_hi = HI64(a), a_lo = LO64(a);
    u128 carry_hi = HI64(*carry), carry_lo = LO64(*carry);
    u128 x = b_lo * c_lo + a_lo + carry_lo;
    u128 y = b_lo * c_hi;
    u128 z = b_hi * c_lo;
    u128 r = (HI64(x) + LO64(y)) + (LO64(z) + a_hi) + carry_hi;
    *carry = (b_hi * c_hi) + HI64(r) + HI64(y) + HI64(z);
    return (LO64(r) << 64) | LO64(x);
}

/* src/arith.rs:473-522: acc[from..] += b * c (one radix-2^128 digit), early-out on c == 0. */
static inline void mac_digit(int from, u128 acc[4], const u128 b[2], u128 c) {
    if (c == 0) return;
    u128 carry = 0;
    for (int i = 0; i < 2; i++) acc[i + from] = mac_with_carry(acc[i + from], b[i], c, &carry);
    for (int i = 0; i < 2; i++) {
        int k = i + from + 2;
        if (k < 4) {
            u128 s = acc[k] + carry;
            carry = s < acc[k];
            acc[k] = s;
        }
    }
}

/* src/arith.rs:525-545 mul_reduce + :300-306 U256::mul.  a <- a*b*R^-1 mod m, fully reduced. */
static inline void u256_mont_mul(u256 *a, const u256 *b, const u256 *m, u128 inv) {
    u128 res[4] = {0, 0, 0, 0};
    for (int i = 0; i < 2; i++) mac_digit(i, res, b->l, a->l[i]);
    for (int i = 0; i < 2; i++) {
        u128 k = inv * res[i];
        mac_digit(i, res, m->l, k);
    }
    a->l[0] = res[2];
    a->l[1] = res[3];
    if (u256_cmp(a, m) >= 0) sub_noborrow(a, m);
}
static inline void u256_mod_add(u256 *a, const u256 *b, const u256 *m) {   /* src/arith.rs:281-287 */
    add_nocarry(a, b);
    if (u256_cmp(a, m) >= 0) sub_noborrow(a, m);
}
static inline void u256_mod_sub(u256 *a, const u256 *b, const u256 *m) {   /* src/arith.rs:290-296 */
    if (u256_cmp(a, b) < 0) add_nocarry(a, m);
    sub_noborrow(a, b);
}
static inline void u256_mod_neg(u256 *a, const u256 *m) {                  /* src/arith.rs:309-316 */
    if (!u256_is_zero(a)) { u256 t = *m; sub_noborrow(&t, a); *a = t; }
}
/* src/arith.rs:324-370: binary extended Euclid (Guajardo et al., Alg. 16). */
static void u256_invert(u256 *a, const u256 *m) {
    u256 u = *a, v = *m, b = {{1, 0}}, c = {{0, 0}}, one = {{1, 0}};
    while (!u256_eq(&u, &one) && !u256_eq(&v, &one)) {
        while ((u.l[0] & 1) == 0) {
            div2(&u);
            if ((b.l[0] & 1) == 0) div2(&b); else { add_nocarry(&b, m); div2(&b); }
        }
        while ((v.l[0] & 1) == 0) {
            div2(&v);
            if ((c.l[0] & 1) == 0) div2(&c); else { add_nocarry(&c, m); div2(&c); }
        }
        if (u256_cmp(&u, &v) >= 0) { sub_noborrow(&u, &v); u256_mod_sub(&b, &c, m); }
        else { sub_noborrow(&v, &u); u256_mod_sub(&c, &b, m); }
    }
    *a = u256_eq(&u, &one) ? b : c;
}
static inline int u256_bit(const u256 *a, int n) { return (int)((a->l[n >> 7] >> (n & 127)) & 1); }

/* ------------------------------------------------------------------------------------------ */
/* Fq / Fr (src/fields/fp.rs:7-222).  Constants derived at init.                                */
/* ------------------------------------------------------------------------------------------ */
typedef u256 fq;
typedef struct { u256 m; u128 inv; u256 one, r2, r3; } fpctx;
static fpctx FQ, FR;
static int g_ready = 0;

static u128 neg_inv128(u128 m0) {               /* -m^-1 mod 2^128 by Newton iteration */
    u128 x = 1;
    for (int i = 0; i < 8; i++) x *= 2 - m0 * x;
    return (u128)0 - x;
}
static void fpctx_init(fpctx *c, const uint64_t w[4]) {
    c->m.l[0] = ((u128)w[1] << 64) | w[0];
    c->m.l[1] = ((u128)w[3] << 64) | w[2];
    c->inv = neg_inv128(c->m.l[0]);
    u256 x = {{1, 0}};                          /* 2^k mod m by repeated doubling */
    for (int i = 0; i < 256; i++) u256_mod_add(&x, &x, &c->m);
    c->one = x;                                 /* R mod m */
    for (int i = 0; i < 256; i++) u256_mod_add(&x, &x, &c->m);
    c->r2 = x;                                  /* R^2 mod m */
    c->r3 = c->r2;
    u256_mont_mul(&c->r3, &c->r2, &c->m, c->inv); /* R^2*R^2/R = R^3 */
}

static inline fq fq_add(fq a, fq b) { u256_mod_add(&a, &b, &FQ.m); return a; }
static inline fq fq_sub(fq a, fq b) { u256_mod_sub(&a, &b, &FQ.m); return a; }
static inline fq fq_neg(fq a) { u256_mod_neg(&a, &FQ.m); return a; }
static inline fq fq_mul(fq a, fq b) { u256_mont_mul(&a, &b, &FQ.m, FQ.inv); return a; }
static inline fq fq_sqr(fq a) { return fq_mul(a, a); }            /* FieldElement::squared default, src/fields/mod.rs:31-33 */
static inline int fq_is_zero(fq a) { return u256_is_zero(&a); }
static inline int fq_eq(fq a, fq b) { return u256_eq(&a, &b); }
static const fq FQ_ZERO = {{0, 0}};
static int fq_inverse(fq a, fq *out) {         /* src/fields/fp.rs:108-117 */
    if (fq_is_zero(a)) return 0;
    u256_invert(&a, &FQ.m);
    u256_mont_mul(&a, &FQ.r3, &FQ.m, FQ.inv);
    *out = a;
    return 1;
}
static fq fq_from_u64(uint64_t v) {            /* Fq::new, src/fields/fp.rs:46-54 */
    fq a = {{v, 0}};
    u256_mont_mul(&a, &FQ.r2, &FQ.m, FQ.inv);
    return a;
}
static fq fq_pow_u256(fq a, const u256 *e) {   /* FieldElement::pow, src/fields/mod.rs:35-46 */
    fq res = FQ.one;
    for (int i = 255; i >= 0; i--) { res = fq_sqr(res); if (u256_bit(e, i)) res = fq_mul(a, res); }
    return res;
}

/* ------------------------------------------------------------------------------------------ */
/* Fq2 (src/fields/fq2.rs)                                                                      */
/* ------------------------------------------------------------------------------------------ */
typedef struct { fq c0, c1; } fq2;
static fq FQ_NONRES;      /* q-1 in Montgomery form, src/fields/fq2.rs:6-16 */
static fq2 XI;            /* 9+u, src/fields/fq2.rs:18-34 */
static fq2 FQ2_ZERO, FQ2_ONE;

static inline fq2 fq2_add(fq2 a, fq2 b) { fq2 r = {fq_add(a.c0, b.c0), fq_add(a.c1, b.c1)}; return r; }
static inline fq2 fq2_sub(fq2 a, fq2 b) { fq2 r = {fq_sub(a.c0, b.c0), fq_sub(a.c1, b.c1)}; return r; }
static inline fq2 fq2_neg(fq2 a) { fq2 r = {fq_neg(a.c0), fq_neg(a.c1)}; return r; }
static inline int fq2_is_zero(fq2 a) { return fq_is_zero(a.c0) && fq_is_zero(a.c1); }
static inline int fq2_eq(fq2 a, fq2 b) { return fq_eq(a.c0, b.c0) && fq_eq(a.c1, b.c1); }
static inline fq2 fq2_mul(fq2 a, fq2 b) {      /* src/fields/fq2.rs:133-149 */
    fq aa = fq_mul(a.c0, b.c0);
    fq bb = fq_mul(a.c1, b.c1);
    fq2 r;
    r.c0 = fq_add(fq_mul(bb, FQ_NONRES), aa);
    r.c1 = fq_sub(fq_sub(fq_mul(fq_add(a.c0, a.c1), fq_add(b.c0, b.c1)), aa), bb);
    return r;
}
static inline fq2 fq2_sqr(fq2 a) {             /* src/fields/fq2.rs:105-117 */
    fq ab = fq_mul(a.c0, a.c1);
    fq2 r;
    r.c0 = fq_sub(fq_sub(fq_mul(fq_add(fq_mul(a.c1, FQ_NONRES), a.c0), fq_add(a.c0, a.c1)), ab),
                  fq_mul(ab, FQ_NONRES));
    r.c1 = fq_add(ab, ab);
    return r;
}
static inline fq2 fq2_scale(fq2 a, fq k) { fq2 r = {fq_mul(a.c0, k), fq_mul(a.c1, k)}; return r; }   /* :48-53 */
static inline fq2 fq2_mul_by_nonresidue(fq2 a) { return fq2_mul(a, XI); }                          /* :55-57 */
static inline fq2 fq2_frobenius(fq2 a, int power) {                                                 /* :59-68 */
    if (power % 2 == 0) return a;
    fq2 r = {a.c0, fq_mul(a.c1, FQ_NONRES)};
    return r;
}
static int fq2_inverse(fq2 a, fq2 *out) {      /* src/fields/fq2.rs:119-130 */
    fq t;
    if (!fq_inverse(fq_sub(fq_sqr(a.c0), fq_mul(fq_sqr(a.c1), FQ_NONRES)), &t)) return 0;
    out->c0 = fq_mul(a.c0, t);
    out->c1 = fq_neg(fq_mul(a.c1, t));
    return 1;
}
static fq2 fq2_pow_u256(fq2 a, const u256 *e) {
    fq2 res = FQ2_ONE;
    for (int i = 255; i >= 0; i--) { res = fq2_sqr(res); if (u256_bit(e, i)) res = fq2_mul(a, res); }
    return res;
}

/* ------------------------------------------------------------------------------------------ */
/* Fq6 (src/fields/fq6.rs)                                                                      */
/* ------------------------------------------------------------------------------------------ */
typedef struct { fq2 c0, c1, c2; } fq6;
static fq2 FROB6_C1[4], FROB6_C2[4], FROB12_C1[4];   /* src/fields/fq6.rs:5-90, src/fields/fq12.rs:6-48 */
static fq6 FQ6_ZERO, FQ6_ONE;

static inline fq6 fq6_add(fq6 a, fq6 b) { fq6 r = {fq2_add(a.c0, b.c0), fq2_add(a.c1, b.c1), fq2_add(a.c2, b.c2)}; return r; }
static inline fq6 fq6_sub(fq6 a, fq6 b) { fq6 r = {fq2_sub(a.c0, b.c0), fq2_sub(a.c1, b.c1), fq2_sub(a.c2, b.c2)}; return r; }
static inline fq6 fq6_neg(fq6 a) { fq6 r = {fq2_neg(a.c0), fq2_neg(a.c1), fq2_neg(a.c2)}; return r; }
static inline fq6 fq6_mul_by_nonresidue(fq6 a) { fq6 r = {fq2_mul_by_nonresidue(a.c2), a.c0, a.c1}; return r; }  /* :109-115 */
static inline fq6 fq6_scale(fq6 a, fq2 k) { fq6 r = {fq2_mul(a.c0, k), fq2_mul(a.c1, k), fq2_mul(a.c2, k)}; return r; } /* :117-123 */
static inline fq6 fq6_frobenius(fq6 a, int p) {                                                                     /* :125-131 */
    fq6 r = {fq2_frobenius(a.c0, p), fq2_mul(fq2_frobenius(a.c1, p), FROB6_C1[p]), fq2_mul(fq2_frobenius(a.c2, p), FROB6_C2[p])};
    return r;
}
static fq6 fq6_mul(fq6 a, fq6 b) {             /* src/fields/fq6.rs:194-208 */
    fq2 a_a = fq2_mul(a.c0, b.c0), b_b = fq2_mul(a.c1, b.c1), c_c = fq2_mul(a.c2, b.c2);
    fq6 r;
    r.c0 = fq2_add(fq2_mul_by_nonresidue(fq2_sub(fq2_sub(fq2_mul(fq2_add(a.c1, a.c2), fq2_add(b.c1, b.c2)), b_b), c_c)), a_a);
    r.c1 = fq2_add(fq2_sub(fq2_sub(fq2_mul(fq2_add(a.c0, a.c1), fq2_add(b.c0, b.c1)), a_a), b_b), fq2_mul_by_nonresidue(c_c));
    r.c2 = fq2_sub(fq2_add(fq2_sub(fq2_mul(fq2_add(a.c0, a.c2), fq2_add(b.c0, b.c2)), a_a), b_b), c_c);
    return r;
}
static fq6 fq6_sqr(fq6 a) {                    /* src/fields/fq6.rs:163-177 */
    fq2 s0 = fq2_sqr(a.c0);
    fq2 ab = fq2_mul(a.c0, a.c1);
    fq2 s1 = fq2_add(ab, ab);
    fq2 s2 = fq2_sqr(fq2_add(fq2_sub(a.c0, a.c1), a.c2));
    fq2 bc = fq2_mul(a.c1, a.c2);
    fq2 s3 = fq2_add(bc, bc);
    fq2 s4 = fq2_sqr(a.c2);
    fq6 r;
    r.c0 = fq2_add(s0, fq2_mul_by_nonresidue(s3));
    r.c1 = fq2_add(s1, fq2_mul_by_nonresidue(s4));
    r.c2 = fq2_sub(fq2_sub(fq2_add(fq2_add(s1, s2), s3), s0), s4);
    return r;
}
static int fq6_inverse(fq6 a, fq6 *out) {      /* src/fields/fq6.rs:179-191 */
    fq2 c0 = fq2_sub(fq2_sqr(a.c0), fq2_mul(a.c1, fq2_mul_by_nonresidue(a.c2)));
    fq2 c1 = fq2_sub(fq2_mul_by_nonresidue(fq2_sqr(a.c2)), fq2_mul(a.c0, a.c1));
    fq2 c2 = fq2_sub(fq2_sqr(a.c1), fq2_mul(a.c0, a.c2));
    fq2 t;
    if (!fq2_inverse(fq2_add(fq2_mul_by_nonresidue(fq2_add(fq2_mul(a.c2, c1), fq2_mul(a.c1, c2))), fq2_mul(a.c0, c0)), &t)) return 0;
    out->c0 = fq2_mul(t, c0); out->c1 = fq2_mul(t, c1); out->c2 = fq2_mul(t, c2);
    return 1;
}

/* ------------------------------------------------------------------------------------------ */
/* Fq12 (src/fields/fq12.rs)                                                                    */
/* ------------------------------------------------------------------------------------------ */
typedef struct { fq6 c0, c1; } fq12;
static fq12 FQ12_ONE;
static u256 BN_U;         /* src/fields/fq12.rs:122 */

static inline fq12 fq12_add(fq12 a, fq12 b) { fq12 r = {fq6_add(a.c0, b.c0), fq6_add(a.c1, b.c1)}; return r; }
static inline fq12 fq12_sub(fq12 a, fq12 b) { fq12 r = {fq6_sub(a.c0, b.c0), fq6_sub(a.c1, b.c1)}; return r; }
static inline fq12 fq12_neg(fq12 a) { fq12 r = {fq6_neg(a.c0), fq6_neg(a.c1)}; return r; }
static fq12 fq12_mul(fq12 a, fq12 b) {         /* src/fields/fq12.rs:316-328 */
    fq6 aa = fq6_mul(a.c0, b.c0), bb = fq6_mul(a.c1, b.c1);
    fq12 r;
    r.c0 = fq6_add(fq6_mul_by_nonresidue(bb), aa);
    r.c1 = fq6_sub(fq6_sub(fq6_mul(fq6_add(a.c0, a.c1), fq6_add(b.c0, b.c1)), aa), bb);
    return r;
}
static fq12 fq12_sqr(fq12 a) {                 /* src/fields/fq12.rs:295-303 */
    fq6 ab = fq6_mul(a.c0, a.c1);
    fq12 r;
    r.c0 = fq6_sub(fq6_sub(fq6_mul(fq6_add(fq6_mul_by_nonresidue(a.c1), a.c0), fq6_add(a.c0, a.c1)), ab), fq6_mul_by_nonresidue(ab));
    r.c1 = fq6_add(ab, ab);
    return r;
}
static int fq12_inverse(fq12 a, fq12 *out) {   /* src/fields/fq12.rs:305-313 */
    fq6 t;
    if (!fq6_inverse(fq6_sub(fq6_sqr(a.c0), fq6_mul_by_nonresidue(fq6_sqr(a.c1))), &t)) return 0;
    out->c0 = fq6_mul(a.c0, t);
    out->c1 = fq6_neg(fq6_mul(a.c1, t));
    return 1;
}
static inline fq12 fq12_unitary_inverse(fq12 a) { fq12 r = {a.c0, fq6_neg(a.c1)}; return r; }   /* :126-128 */
static fq12 fq12_frobenius(fq12 a, int p) {     /* src/fields/fq12.rs:112-119 */
    fq12 r = {fq6_frobenius(a.c0, p), fq6_scale(fq6_frobenius(a.c1, p), FROB12_C1[p])};
    return r;
}
/* src/fields/fq12.rs:130-196, statement for statement */
static fq12 fq12_mul_by_024(fq12 f, fq2 ell_0, fq2 ell_vw, fq2 ell_vv) {
    fq2 z0 = f.c0.c0, z1 = f.c0.c1, z2 = f.c0.c2, z3 = f.c1.c0, z4 = f.c1.c1, z5 = f.c1.c2;
    fq2 x0 = ell_0, x2 = ell_vv, x4 = ell_vw;
    fq2 d0 = fq2_mul(z0, x0), d2 = fq2_mul(z2, x2), d4 = fq2_mul(z4, x4);
    fq2 t2 = fq2_add(z0, z4), t1 = fq2_add(z0, z2), s0 = fq2_add(fq2_add(z1, z3), z5);
    fq2 s1 = fq2_mul(z1, x2);
    fq2 t3 = fq2_add(s1, d4);
    fq2 t4 = fq2_add(fq2_mul_by_nonresidue(t3), d0);
    z0 = t4;
    t3 = fq2_mul(z5, x4); s1 = fq2_add(s1, t3); t3 = fq2_add(t3, d2); t4 = fq2_mul_by_nonresidue(t3);
    t3 = fq2_mul(z1, x0); s1 = fq2_add(s1, t3); t4 = fq2_add(t4, t3);
    z1 = t4;
    fq2 t0 = fq2_add(x0, x2);
    t3 = fq2_sub(fq2_sub(fq2_mul(t1, t0), d0), d2);
    t4 = fq2_mul(z3, x4); s1 = fq2_add(s1, t4); t3 = fq2_add(t3, t4);
    t0 = fq2_add(z2, z4);
    z2 = t3;
    t1 = fq2_add(x2, x4);
    t3 = fq2_sub(fq2_sub(fq2_mul(t0, t1), d2), d4);
    t4 = fq2_mul_by_nonresidue(t3);
    t3 = fq2_mul(z3, x0); s1 = fq2_add(s1, t3); t4 = fq2_add(t4, t3);
    z3 = t4;
    t3 = fq2_mul(z5, x2); s1 = fq2_add(s1, t3); t4 = fq2_mul_by_nonresidue(t3);
    t0 = fq2_add(x0, x4);
    t3 = fq2_sub(fq2_sub(fq2_mul(t2, t0), d0), d4);
    t4 = fq2_add(t4, t3);
    z4 = t4;
    t0 = fq2_add(fq2_add(x0, x2), x4);
    t3 = fq2_sub(fq2_mul(s0, t0), s1);
    z5 = t3;
    fq12 r = {{z0, z1, z2}, {z3, z4, z5}};
    return r;
}
/* src/fields/fq12.rs:198-247 */
static fq12 fq12_cyclotomic_squared(fq12 a) {
    fq2 z0 = a.c0.c0, z4 = a.c0.c1, z3 = a.c0.c2, z2 = a.c1.c0, z1 = a.c1.c1, z5 = a.c1.c2;
    fq2 tmp = fq2_mul(z0, z1);
    fq2 t0 = fq2_sub(fq2_sub(fq2_mul(fq2_add(z0, z1), fq2_add(fq2_mul_by_nonresidue(z1), z0)), tmp), fq2_mul_by_nonresidue(tmp));
    fq2 t1 = fq2_add(tmp, tmp);
    tmp = fq2_mul(z2, z3);
    fq2 t2 = fq2_sub(fq2_sub(fq2_mul(fq2_add(z2, z3), fq2_add(fq2_mul_by_nonresidue(z3), z2)), tmp), fq2_mul_by_nonresidue(tmp));
    fq2 t3 = fq2_add(tmp, tmp);
    tmp = fq2_mul(z4, z5);
    fq2 t4 = fq2_sub(fq2_sub(fq2_mul(fq2_add(z4, z5), fq2_add(fq2_mul_by_nonresidue(z5), z4)), tmp), fq2_mul_by_nonresidue(tmp));
    fq2 t5 = fq2_add(tmp, tmp);
    z0 = fq2_sub(t0, z0); z0 = fq2_add(z0, z0); z0 = fq2_add(z0, t0);
    z1 = fq2_add(t1, z1); z1 = fq2_add(z1, z1); z1 = fq2_add(z1, t1);
    tmp = fq2_mul_by_nonresidue(t5);
    z2 = fq2_add(tmp, z2); z2 = fq2_add(z2, z2); z2 = fq2_add(z2, tmp);
    z3 = fq2_sub(t4, z3); z3 = fq2_add(z3, z3); z3 = fq2_add(z3, t4);
    z4 = fq2_sub(t2, z4); z4 = fq2_add(z4, z4); z4 = fq2_add(z4, t2);
    z5 = fq2_add(t3, z5); z5 = fq2_add(z5, z5); z5 = fq2_add(z5, t3);
    fq12 r = {{z0, z4, z3}, {z2, z1, z5}};
    return r;
}
static fq12 fq12_cyclotomic_pow(fq12 a, const u256 *e) {   /* src/fields/fq12.rs:249-266 */
    fq12 res = FQ12_ONE;
    int found_one = 0;
    for (int i = 255; i >= 0; i--) {
        if (found_one) res = fq12_cyclotomic_squared(res);
        if (u256_bit(e, i)) { found_one = 1; res = fq12_mul(a, res); }
    }
    return res;
}
static fq12 fq12_exp_by_neg_z(fq12 a) { return fq12_unitary_inverse(fq12_cyclotomic_pow(a, &BN_U)); }  /* :121-124 */
static int fq12_final_exp_first_chunk(fq12 a, fq12 *out) {   /* src/fields/fq12.rs:62-73 */
    fq12 b;
    if (!fq12_inverse(a, &b)) return 0;
    fq12 c = fq12_mul(fq12_unitary_inverse(a), b);
    *out = fq12_mul(fq12_frobenius(c, 2), c);
    return 1;
}
static fq12 fq12_final_exp_last_chunk(fq12 x) {              /* src/fields/fq12.rs:75-105 */
    fq12 a = fq12_exp_by_neg_z(x);
    fq12 b = fq12_cyclotomic_squared(a);
    fq12 c = fq12_cyclotomic_squared(b);
    fq12 d = fq12_mul(c, b);
    fq12 e = fq12_exp_by_neg_z(d);
    fq12 f = fq12_cyclotomic_squared(e);
    fq12 g = fq12_exp_by_neg_z(f);
    fq12 h = fq12_unitary_inverse(d);
    fq12 i = fq12_unitary_inverse(g);
    fq12 j = fq12_mul(i, e);
    fq12 k = fq12_mul(j, h);
    fq12 l = fq12_mul(k, b);
    fq12 m = fq12_mul(k, e);
    fq12 n = fq12_mul(x, m);
    fq12 o = fq12_frobenius(l, 1);
    fq12 p = fq12_mul(o, n);
    fq12 q = fq12_frobenius(k, 2);
    fq12 r = fq12_mul(q, p);
    fq12 s = fq12_unitary_inverse(x);
    fq12 t = fq12_mul(s, l);
    fq12 u = fq12_frobenius(t, 3);
    return fq12_mul(u, r);
}
static int fq12_final_exponentiation(fq12 a, fq12 *out) {    /* src/fields/fq12.rs:107-110 */
    fq12 fc;
    if (!fq12_final_exp_first_chunk(a, &fc)) return 0;
    *out = fq12_final_exp_last_chunk(fc);
    return 1;
}
static fq12 fq12_pow_u256(fq12 a, const u256 *e) {           /* src/fields/mod.rs:35-46 (Gt::pow) */
    fq12 res = FQ12_ONE;
    for (int i = 255; i >= 0; i--) { res = fq12_sqr(res); if (u256_bit(e, i)) res = fq12_mul(a, res); }
    return res;
}

/* ------------------------------------------------------------------------------------------ */
/* Groups (src/groups/mod.rs:198-369)                                                           */
/* ------------------------------------------------------------------------------------------ */
typedef struct { fq x, y, z; } g1;
typedef struct { fq2 x, y, z; } g2;
typedef struct { fq x, y; } g1a;
typedef struct { fq2 x, y; } g2a;

#define DEFINE_GROUP(G, F, FADD, FSUB, FNEG, FMUL, FSQR, FISZERO, FEQ, FZERO, FONE)                                   \
    static G G##_zero(void) { G r = {FZERO, FONE, FZERO}; return r; }                        /* :230-236 */        \
    static G G##_double(G p) {                                                               /* :250-269 */        \
        F a = FSQR(p.x), b = FSQR(p.y), c = FSQR(b);                                                                  \
        F d = FSUB(FSUB(FSQR(FADD(p.x, b)), a), c);                                                                   \
        d = FADD(d, d);                                                                                               \
        F e = FADD(FADD(a, a), a);                                                                                    \
        F f = FSQR(e);                                                                                                \
        F x3 = FSUB(f, FADD(d, d));                                                                                   \
        F eight_c = FADD(c, c);                                                                                       \
        eight_c = FADD(eight_c, eight_c);                                                                             \
        eight_c = FADD(eight_c, eight_c);                                                                             \
        F y1z1 = FMUL(p.y, p.z);                                                                                      \
        G r = {x3, FSUB(FMUL(e, FSUB(d, x3)), eight_c), FADD(y1z1, y1z1)};                                            \
        return r;                                                                                                     \
    }                                                                                                                 \
    static G G##_add(G p, G o) {                                                             /* :294-334 */        \
        if (FISZERO(p.z)) return o;                                                                                   \
        if (FISZERO(o.z)) return p;                                                                                   \
        F z1_squared = FSQR(p.z), z2_squared = FSQR(o.z);                                                             \
        F u1 = FMUL(p.x, z2_squared), u2 = FMUL(o.x, z1_squared);                                                     \
        F z1_cubed = FMUL(p.z, z1_squared), z2_cubed = FMUL(o.z, z2_squared);                                         \
        F s1 = FMUL(p.y, z2_cubed), s2 = FMUL(o.y, z1_cubed);                                                         \
        if (FEQ(u1, u2) && FEQ(s1, s2)) return G##_double(p);                                                         \
        F h = FSUB(u2, u1), s2_minus_s1 = FSUB(s2, s1);                                                               \
        F i = FSQR(FADD(h, h));                                                                                       \
        F j = FMUL(h, i);                                                                                             \
        F r = FADD(s2_minus_s1, s2_minus_s1);                                                                         \
        F v = FMUL(u1, i);                                                                                            \
        F s1_j = FMUL(s1, j);                                                                                         \
        F x3 = FSUB(FSUB(FSQR(r), j), FADD(v, v));                                                                    \
        G res = {x3, FSUB(FMUL(r, FSUB(v, x3)), FADD(s1_j, s1_j)),                                                    \
                 FMUL(FSUB(FSUB(FSQR(FADD(p.z, o.z)), z1_squared), z2_squared), h)};                                  \
        return res;                                                                                                   \
    }                                                                                                                 \
    static G G##_mul(G p, const u256 *k) {                      /* :272-292, k = canonical scalar bits */            \
        G res = G##_zero();                                                                                           \
        int found_one = 0;                                                                                            \
        for (int i = 255; i >= 0; i--) {                                                                              \
            if (found_one) res = G##_double(res);                                                                     \
            if (u256_bit(k, i)) { found_one = 1; res = G##_add(res, p); }                                             \
        }                                                                                                             \
        return res;                                                                                                   \
    }

DEFINE_GROUP(g1, fq, fq_add, fq_sub, fq_neg, fq_mul, fq_sqr, fq_is_zero, fq_eq, FQ_ZERO, FQ.one)
DEFINE_GROUP(g2, fq2, fq2_add, fq2_sub, fq2_neg, fq2_mul, fq2_sqr, fq2_is_zero, fq2_eq, FQ2_ZERO, FQ2_ONE)

static int g1_to_affine(g1 p, g1a *out) {      /* src/groups/mod.rs:198-217 */
    if (fq_is_zero(p.z)) return 0;
    if (fq_eq(p.z, FQ.one)) { out->x = p.x; out->y = p.y; return 1; }
    fq zinv; fq_inverse(p.z, &zinv);
    fq zinv2 = fq_sqr(zinv);
    out->x = fq_mul(p.x, zinv2);
    out->y = fq_mul(p.y, fq_mul(zinv2, zinv));
    return 1;
}
static int g2_to_affine(g2 p, g2a *out) {
    if (fq2_is_zero(p.z)) return 0;
    if (fq2_eq(p.z, FQ2_ONE)) { out->x = p.x; out->y = p.y; return 1; }
    fq2 zinv; fq2_inverse(p.z, &zinv);
    fq2 zinv2 = fq2_sqr(zinv);
    out->x = fq2_mul(p.x, zinv2);
    out->y = fq2_mul(p.y, fq2_mul(zinv2, zinv));
    return 1;
}

/* ------------------------------------------------------------------------------------------ */
/* Pairing (src/groups/mod.rs:515-926)                                                          */
/* ------------------------------------------------------------------------------------------ */
static const uint8_t NAF[64] = {1,0,1,0,0,0,3,0,3,0,0,0,3,0,1,0,3,0,0,3,0,0,0,0,0,1,0,0,3,0,1,0,
                                0,3,0,0,0,0,3,0,1,0,0,0,3,0,3,0,0,1,0,0,0,3,0,0,3,0,1,0,1,0,0,0};  /* :9-14 */
#define N_COEFFS 87
typedef struct { fq2 ell_0, ell_vw, ell_vv; } ellcoeffs;    /* :566-571, field order of the struct */
typedef struct { g2a q; ellcoeffs coeffs[N_COEFFS]; } g2precomp;   /* :573-577 (Vec flattened) */
static fq TWO_INV;                 /* :520-528 */
static fq2 G2_B;                   /* :452-467 */
static fq2 TWIST_MUL_BY_Q_X, TWIST_MUL_BY_Q_Y;   /* :530-564 */
static g1 G1_ONE;
static g2 G2_ONE;

static ellcoeffs g2_doubling_step(g2 *r) {     /* :754-776 */
    fq2 a = fq2_scale(fq2_mul(r->x, r->y), TWO_INV);
    fq2 b = fq2_sqr(r->y);
    fq2 c = fq2_sqr(r->z);
    fq2 d = fq2_add(fq2_add(c, c), c);
    fq2 e = fq2_mul(G2_B, d);
    fq2 f = fq2_add(fq2_add(e, e), e);
    fq2 g = fq2_scale(fq2_add(b, f), TWO_INV);
    fq2 h = fq2_sub(fq2_sqr(fq2_add(r->y, r->z)), fq2_add(b, c));
    fq2 i = fq2_sub(e, b);
    fq2 j = fq2_sqr(r->x);
    fq2 e_sq = fq2_sqr(e);
    r->x = fq2_mul(a, fq2_sub(b, f));
    r->y = fq2_sub(fq2_sqr(g), fq2_add(fq2_add(e_sq, e_sq), e_sq));
    r->z = fq2_mul(b, h);
    ellcoeffs out = {fq2_mul(XI, i), fq2_neg(h), fq2_add(fq2_add(j, j), j)};
    return out;
}
static ellcoeffs g2_mixed_addition_step(g2 *r, const g2a *base) {   /* :731-752 */
    fq2 d = fq2_sub(r->x, fq2_mul(r->z, base->x));
    fq2 e = fq2_sub(r->y, fq2_mul(r->z, base->y));
    fq2 f = fq2_sqr(d);
    fq2 g = fq2_sqr(e);
    fq2 h = fq2_mul(d, f);
    fq2 i = fq2_mul(r->x, f);
    fq2 j = fq2_sub(fq2_add(fq2_mul(r->z, g), h), fq2_add(i, i));
    r->x = fq2_mul(d, j);
    r->y = fq2_sub(fq2_mul(e, fq2_sub(i, j)), fq2_mul(h, r->y));
    r->z = fq2_mul(r->z, h);
    ellcoeffs out = {fq2_mul(XI, fq2_sub(fq2_mul(e, base->x), fq2_mul(d, base->y))), d, fq2_neg(e)};
    return out;
}
static g2a g2a_mul_by_q(g2a q) {               /* :694-699 */
    g2a r = {fq2_mul(TWIST_MUL_BY_Q_X, fq2_frobenius(q.x, 1)), fq2_mul(TWIST_MUL_BY_Q_Y, fq2_frobenius(q.y, 1))};
    return r;
}
static void g2_precompute(g2a q, g2precomp *out) {   /* :701-728 */
    g2 r = {q.x, q.y, FQ2_ONE};
    int n = 0;
    g2a q_neg = {q.x, fq2_neg(q.y)};
    for (int k = 0; k < 64; k++) {
        out->coeffs[n++] = g2_doubling_step(&r);
        if (NAF[k] == 1) out->coeffs[n++] = g2_mixed_addition_step(&r, &q);
        if (NAF[k] == 3) out->coeffs[n++] = g2_mixed_addition_step(&r, &q_neg);
    }
    g2a q1 = g2a_mul_by_q(q);
    g2a q2 = g2a_mul_by_q(q1);
    q2.y = fq2_neg(q2.y);
    out->coeffs[n++] = g2_mixed_addition_step(&r, &q1);
    out->coeffs[n++] = g2_mixed_addition_step(&r, &q2);
    out->q = q;
}
static inline fq12 line_mul(fq12 f, const ellcoeffs *c, const g1a *p) {     /* the call shape at :589 */
    return fq12_mul_by_024(f, c->ell_0, fq2_scale(c->ell_vw, p->y), fq2_scale(c->ell_vv, p->x));
}
static fq12 miller_loop_batch(const g2precomp *pcs, const g1a *ps, size_t n) {   /* :609-640; n==1 is :580-606 */
    fq12 f = FQ12_ONE;
    int idx = 0;
    for (int k = 0; k < 64; k++) {
        f = fq12_sqr(f);
        for (size_t j = 0; j < n; j++) f = line_mul(f, &pcs[j].coeffs[idx], &ps[j]);
        idx++;
        if (NAF[k] != 0) {
            for (size_t j = 0; j < n; j++) f = line_mul(f, &pcs[j].coeffs[idx], &ps[j]);
            idx++;
        }
    }
    for (size_t j = 0; j < n; j++) f = line_mul(f, &pcs[j].coeffs[idx], &ps[j]);
    idx++;
    for (size_t j = 0; j < n; j++) f = line_mul(f, &pcs[j].coeffs[idx], &ps[j]);
    return f;
}
static fq12 pairing(g1 p, g2 q) {              /* :894-902 */
    g1a pa; g2a qa;
    if (!g1_to_affine(p, &pa) || !g2_to_affine(q, &qa)) return FQ12_ONE;
    g2precomp pc;
    g2_precompute(qa, &pc);
    fq12 out = FQ12_ONE;
    fq12_final_exponentiation(miller_loop_batch(&pc, &pa, 1), &out);
    return out;
}

/* ------------------------------------------------------------------------------------------ */
/* init: derive every constant from q, r, xi, u                                                 */
/* ------------------------------------------------------------------------------------------ */
static void u256_div_small(u256 *a, uint64_t d) {   /* exact schoolbook division by a small integer */
    uint64_t w[4] = {(uint64_t)a->l[0], (uint64_t)(a->l[0] >> 64), (uint64_t)a->l[1], (uint64_t)(a->l[1] >> 64)};
    u128 rem = 0;
    for (int i = 3; i >= 0; i--) { u128 cur = (rem << 64) | w[i]; w[i] = (uint64_t)(cur / d); rem = cur % d; }
    a->l[0] = ((u128)w[1] << 64) | w[0];
    a->l[1] = ((u128)w[3] << 64) | w[2];
}
static fq2 fq2_conj(fq2 a) { fq2 r = {a.c0, fq_neg(a.c1)}; return r; }

void orc_init(void) {
    if (g_ready) return;
    static const uint64_t QW[4] = {0x3c208c16d87cfd47ULL, 0x97816a916871ca8dULL, 0xb85045b68181585dULL, 0x30644e72e131a029ULL};
    static const uint64_t RW[4] = {0x43e1f593f0000001ULL, 0x2833e84879b97091ULL, 0xb85045b68181585dULL, 0x30644e72e131a029ULL};
    fpctx_init(&FQ, QW);
    fpctx_init(&FR, RW);
    FQ_NONRES = fq_neg(FQ.one);
    FQ2_ZERO.c0 = FQ_ZERO; FQ2_ZERO.c1 = FQ_ZERO;
    FQ2_ONE.c0 = FQ.one; FQ2_ONE.c1 = FQ_ZERO;
    XI.c0 = fq_from_u64(9); XI.c1 = FQ.one;
    FQ6_ZERO.c0 = FQ6_ZERO.c1 = FQ6_ZERO.c2 = FQ2_ZERO;
    FQ6_ONE = FQ6_ZERO; FQ6_ONE.c0 = FQ2_ONE;
    FQ12_ONE.c0 = FQ6_ONE; FQ12_ONE.c1 = FQ6_ZERO;
    BN_U.l[0] = 4965661367192848881ULL; BN_U.l[1] = 0;
    fq_inverse(fq_from_u64(2), &TWO_INV);
    fq2 xi_inv; fq2_inverse(XI, &xi_inv);
    G2_B = fq2_scale(xi_inv, fq_from_u64(3));
    /* xi^((q-1)/6), xi^((q-1)/3), xi^((q-1)/2); higher Frobenius powers through x^q = conj(x) on Fq2:
       g^( (q^2-1)/(q-1) ) = g*conj(g),  g^( (q^3-1)/(q-1) ) = g*conj(g)*g. */
    u256 e = FQ.m; e.l[0] -= 1;
    u256 e6 = e, e3 = e, e2 = e;
    u256_div_small(&e6, 6); u256_div_small(&e3, 3); u256_div_small(&e2, 2);
    fq2 g12 = fq2_pow_u256(XI, &e6);
    fq2 g61 = fq2_pow_u256(XI, &e3);
    fq2 g62 = fq2_sqr(g61);
    fq2 *tabs[3] = {FROB12_C1, FROB6_C1, FROB6_C2};
    fq2 gens[3] = {g12, g61, g62};
    for (int t = 0; t < 3; t++) {
        tabs[t][0] = FQ2_ONE;
        tabs[t][1] = gens[t];
        tabs[t][2] = fq2_mul(gens[t], fq2_conj(gens[t]));
        tabs[t][3] = fq2_mul(tabs[t][2], gens[t]);
    }
    TWIST_MUL_BY_Q_X = g61;
    TWIST_MUL_BY_Q_Y = fq2_pow_u256(XI, &e2);
    G1_ONE.x = FQ.one; G1_ONE.y = fq_from_u64(2); G1_ONE.z = FQ.one;
    /* G2 generator (canonical decimal values as in the reference's predefined_pair test,
       src/groups/mod.rs:1010-1021), little-endian u64 words of the canonical integers */
    static const uint64_t GX0[4] = {0x46debd5cd992f6edULL, 0x674322d4f75edaddULL, 0x426a00665e5c4479ULL, 0x1800deef121f1e76ULL};
    static const uint64_t GX1[4] = {0x97e485b7aef312c2ULL, 0xf1aa493335a9e712ULL, 0x7260bfb731fb5d25ULL, 0x198e9393920d483aULL};
    static const uint64_t GY0[4] = {0x4ce6cc0166fa7daaULL, 0xe3d1e7690c43d37bULL, 0x4aab71808dcb408fULL, 0x12c85ea5db8c6debULL};
    static const uint64_t GY1[4] = {0x55acdadcd122975bULL, 0xbc4b313370b38ef3ULL, 0xec9e99ad690c3395ULL, 0x090689d0585ff075ULL};
    const uint64_t *gw[4] = {GX0, GX1, GY0, GY1};
    fq *dst[4] = {&G2_ONE.x.c0, &G2_ONE.x.c1, &G2_ONE.y.c0, &G2_ONE.y.c1};
    for (int i = 0; i < 4; i++) {
        fq a; a.l[0] = ((u128)gw[i][1] << 64) | gw[i][0]; a.l[1] = ((u128)gw[i][3] << 64) | gw[i][2];
        u256_mont_mul(&a, &FQ.r2, &FQ.m, FQ.inv);
        *dst[i] = a;
    }
    G2_ONE.z = FQ2_ONE;
    g_ready = 1;
}

/* ------------------------------------------------------------------------------------------ */
/* exported C entry points (byte buffers = repr(C) Montgomery images, 8-byte alignment assumed) */
/* ------------------------------------------------------------------------------------------ */
#define LD(T, base, i) ({ T _v; memcpy(&_v, (const uint8_t *)(base) + (size_t)(i) * sizeof(T), sizeof(T)); _v; })
#define ST(T, base, i, v) do { T _v = (v); memcpy((uint8_t *)(base) + (size_t)(i) * sizeof(T), &_v, sizeof(T)); } while (0)

static int clamp_threads(int t) {
#ifdef _OPENMP
    int mx = omp_get_num_procs();
    if (t <= 0 || t > mx) t = mx;
    return t;
#else
    (void)t; return 1;
#endif
}
int orc_num_procs(void) {
#ifdef _OPENMP
    return omp_get_num_procs();
#else
    return 1;
#endif
}

/* constants dump for tests: see tests/test_oracle_c.py for the order */
void orc_constants(uint8_t *out) {
    orc_init();
    uint8_t *p = out;
#define PUT(x) do { memcpy(p, &(x), sizeof(x)); p += sizeof(x); } while (0)
    PUT(FQ.one); PUT(FQ.r2); PUT(FQ.r3); PUT(FR.one); PUT(FR.r2); PUT(FR.r3);
    PUT(FQ.inv); PUT(FR.inv);
    PUT(FQ_NONRES); PUT(XI); PUT(TWO_INV); PUT(G2_B); PUT(TWIST_MUL_BY_Q_X); PUT(TWIST_MUL_BY_Q_Y);
    for (int k = 1; k <= 3; k++) PUT(FROB6_C1[k]);
    for (int k = 1; k <= 3; k++) PUT(FROB6_C2[k]);
    for (int k = 1; k <= 3; k++) PUT(FROB12_C1[k]);
    PUT(G1_ONE); PUT(G2_ONE);
#undef PUT
}

/* field-level batched ops (used to check the CUDA field kernels) -------------------------------- */
void orc_fq_mul_many(const void *a, const void *b, void *out, size_t n) { orc_init(); for (size_t i = 0; i < n; i++) ST(fq, out, i, fq_mul(LD(fq, a, i), LD(fq, b, i))); }
void orc_fq_add_many(const void *a, const void *b, void *out, size_t n) { orc_init(); for (size_t i = 0; i < n; i++) ST(fq, out, i, fq_add(LD(fq, a, i), LD(fq, b, i))); }
void orc_fq_sub_many(const void *a, const void *b, void *out, size_t n) { orc_init(); for (size_t i = 0; i < n; i++) ST(fq, out, i, fq_sub(LD(fq, a, i), LD(fq, b, i))); }
void orc_fq_inv_many(const void *a, void *out, uint8_t *some, size_t n) {
    orc_init();
    for (size_t i = 0; i < n; i++) { fq r = FQ_ZERO; some[i] = (uint8_t)fq_inverse(LD(fq, a, i), &r); ST(fq, out, i, r); }
}
void orc_fq2_mul_many(const void *a, const void *b, void *out, size_t n) { orc_init(); for (size_t i = 0; i < n; i++) ST(fq2, out, i, fq2_mul(LD(fq2, a, i), LD(fq2, b, i))); }
void orc_fq2_sqr_many(const void *a, void *out, size_t n) { orc_init(); for (size_t i = 0; i < n; i++) ST(fq2, out, i, fq2_sqr(LD(fq2, a, i))); }
void orc_fq6_mul_many(const void *a, const void *b, void *out, size_t n) { orc_init(); for (size_t i = 0; i < n; i++) ST(fq6, out, i, fq6_mul(LD(fq6, a, i), LD(fq6, b, i))); }
void orc_fq12_mul_many(const void *a, const void *b, void *out, size_t n) { orc_init(); for (size_t i = 0; i < n; i++) ST(fq12, out, i, fq12_mul(LD(fq12, a, i), LD(fq12, b, i))); }
void orc_fq12_sqr_many(const void *a, void *out, size_t n) { orc_init(); for (size_t i = 0; i < n; i++) ST(fq12, out, i, fq12_sqr(LD(fq12, a, i))); }
void orc_fq12_add_many(const void *a, const void *b, void *out, size_t n) { orc_init(); for (size_t i = 0; i < n; i++) ST(fq12, out, i, fq12_add(LD(fq12, a, i), LD(fq12, b, i))); }
void orc_fq12_sub_many(const void *a, const void *b, void *out, size_t n) { orc_init(); for (size_t i = 0; i < n; i++) ST(fq12, out, i, fq12_sub(LD(fq12, a, i), LD(fq12, b, i))); }
void orc_fq12_neg_many(const void *a, void *out, size_t n) { orc_init(); for (size_t i = 0; i < n; i++) ST(fq12, out, i, fq12_neg(LD(fq12, a, i))); }
void orc_fq12_cyclotomic_squared_many(const void *a, void *out, size_t n) { orc_init(); for (size_t i = 0; i < n; i++) ST(fq12, out, i, fq12_cyclotomic_squared(LD(fq12, a, i))); }
void orc_fq12_exp_by_neg_z_many(const void *a, void *out, size_t n) { orc_init(); for (size_t i = 0; i < n; i++) ST(fq12, out, i, fq12_exp_by_neg_z(LD(fq12, a, i))); }
void orc_fq12_frobenius_many(const void *a, int power, void *out, size_t n) { orc_init(); for (size_t i = 0; i < n; i++) ST(fq12, out, i, fq12_frobenius(LD(fq12, a, i), power)); }
void orc_fq12_mul_by_024_many(const void *f, const void *ell /* n x (ell_0, ell_vw, ell_vv) */, void *out, size_t n) {
    orc_init();
    for (size_t i = 0; i < n; i++) { ellcoeffs c = LD(ellcoeffs, ell, i); ST(fq12, out, i, fq12_mul_by_024(LD(fq12, f, i), c.ell_0, c.ell_vw, c.ell_vv)); }
}
void orc_fq12_inverse_many(const void *a, void *out, uint8_t *some, size_t n) {
    orc_init();
    for (size_t i = 0; i < n; i++) { fq12 r; memset(&r, 0, sizeof r); some[i] = (uint8_t)fq12_inverse(LD(fq12, a, i), &r); ST(fq12, out, i, r); }
}

/* Gt API: lib.rs:584-609 ----------------------------------------------------------------------- */
void orc_gt_pow_many(const void *a, const void *fr, void *out, size_t n, int threads) {   /* Gt::pow(Fr), lib.rs:592 */
    orc_init();
    threads = clamp_threads(threads);
#pragma omp parallel for num_threads(threads) schedule(static)
    for (size_t i = 0; i < n; i++) {
        u256 k = LD(u256, fr, i), one = {{1, 0}};
        u256_mont_mul(&k, &one, &FR.m, FR.inv);               /* U256::from(Fr), fp.rs:13-20 */
        ST(fq12, out, i, fq12_pow_u256(LD(fq12, a, i), &k));
    }
}
void orc_final_exp_many(const void *a, void *out, uint8_t *some, size_t n, int threads) {   /* lib.rs:598-600 */
    orc_init();
    threads = clamp_threads(threads);
#pragma omp parallel for num_threads(threads) schedule(static)
    for (size_t i = 0; i < n; i++) { fq12 r; memset(&r, 0, sizeof r); some[i] = (uint8_t)fq12_final_exponentiation(LD(fq12, a, i), &r); ST(fq12, out, i, r); }
}

/* groups ---------------------------------------------------------------------------------------- */
void orc_g1_mul_many(const void *p, const void *fr, void *out, size_t n, int threads) {    /* G1 * Fr, lib.rs:425-431 */
    orc_init();
    threads = clamp_threads(threads);
#pragma omp parallel for num_threads(threads) schedule(static)
    for (size_t i = 0; i < n; i++) {
        u256 k = LD(u256, fr, i), one = {{1, 0}};
        u256_mont_mul(&k, &one, &FR.m, FR.inv);
        ST(g1, out, i, g1_mul(LD(g1, p, i), &k));
    }
}
void orc_g2_mul_many(const void *p, const void *fr, void *out, size_t n, int threads) {    /* G2 * Fr, lib.rs:576-582 */
    orc_init();
    threads = clamp_threads(threads);
#pragma omp parallel for num_threads(threads) schedule(static)
    for (size_t i = 0; i < n; i++) {
        u256 k = LD(u256, fr, i), one = {{1, 0}};
        u256_mont_mul(&k, &one, &FR.m, FR.inv);
        ST(g2, out, i, g2_mul(LD(g2, p, i), &k));
    }
}
/* AffineG::new, src/groups/mod.rs:94-113: 0 = Ok, 1 = Err(NotOnCurve), 2 = Err(NotInSubgroup).  Input: affine (x, y)
 * Montgomery images.  G1Params::check_order() is false (:401), G2Params::check_order() is true (:469-471). */
void orc_g1_affine_check_many(const void *xy, uint8_t *status, size_t n) {
    orc_init();
    fq b = fq_from_u64(3);
    for (size_t i = 0; i < n; i++) {
        fq x = LD(fq, xy, 2 * i), y = LD(fq, xy, 2 * i + 1);
        status[i] = fq_eq(fq_sqr(y), fq_add(fq_mul(fq_sqr(x), x), b)) ? 0 : 1;
    }
}
void orc_g2_affine_check_many(const void *xy, uint8_t *status, size_t n, int threads) {
    orc_init();
    threads = clamp_threads(threads);
    u256 rm1 = FR.m;
    rm1.l[0] -= 1;                       /* r - 1 = canonical value of -Fr::one(); r is odd, no borrow */
#pragma omp parallel for num_threads(threads) schedule(static)
    for (size_t i = 0; i < n; i++) {
        fq2 x = LD(fq2, xy, 2 * i), y = LD(fq2, xy, 2 * i + 1);
        if (!fq2_eq(fq2_sqr(y), fq2_add(fq2_mul(fq2_sqr(x), x), G2_B))) { status[i] = 1; continue; }
        g2 p;
        p.x = x; p.y = y; p.z = FQ2_ONE;
        g2 t = g2_add(g2_mul(p, &rm1), p);
        status[i] = fq2_is_zero(t.z) ? 0 : 2;
    }
}

/* ------------------------------------------------------------------------------------------ */
/* square roots and the compressed wire formats                                   */
/* ------------------------------------------------------------------------------------------ */
static u256 fq_canonical(fq a) {               /* From<Fq> for U256, src/fields/fp.rs:13-20: one Montgomery mul by 1 */
    u256 one = {{1, 0}};
    u256_mont_mul(&a, &one, &FQ.m, FQ.inv);
    return a;
}
static u256 fq_minus3_div4(void) {             /* src/fields/fp.rs:233-236: (-3) * 4^-1 = (q-3)/4 since q = 3 mod 4 */
    u256 e = FQ.m, three = {{3, 0}};
    sub_noborrow(&e, &three); div2(&e); div2(&e);
    return e;
}
static u256 fq_minus1_div2(void) {             /* src/fields/fp.rs:238-241: (q-1)/2 */
    u256 e = FQ.m, one = {{1, 0}};
    sub_noborrow(&e, &one); div2(&e);
    return e;
}
static int fq_sqrt(fq a, fq *out) {            /* Fq::sqrt, src/fields/fp.rs:245-260 */
    u256 e = fq_minus3_div4();
    fq a1 = fq_pow_u256(a, &e);
    fq a1a = fq_mul(a1, a);
    fq a0 = fq_mul(a1, a1a);
    if (fq_eq(a0, fq_neg(FQ.one))) return 0;
    *out = a1a;
    return 1;
}
static int fq2_sqrt(fq2 a, fq2 *out) {         /* Fq2::sqrt, src/fields/fq2.rs:208-224 */
    u256 e34 = fq_minus3_div4(), e12 = fq_minus1_div2();
    fq2 a1 = fq2_pow_u256(a, &e34);
    fq2 a1a = fq2_mul(a1, a);
    fq2 alpha = fq2_mul(a1, a1a);
    fq2 a0 = fq2_mul(fq2_pow_u256(alpha, &FQ.m), alpha);
    fq2 m1 = fq2_neg(FQ2_ONE);
    if (fq2_eq(a0, m1)) return 0;
    if (fq2_eq(alpha, m1)) {
        fq2 i = {FQ_ZERO, FQ.one};
        *out = fq2_mul(i, a1a);
    } else {
        fq2 b = fq2_pow_u256(fq2_add(alpha, FQ2_ONE), &e12);
        *out = fq2_mul(b, a1a);
    }
    return 1;
}
static u128 be128(const uint8_t *s) { u128 v = 0; for (int i = 0; i < 16; i++) v = (v << 8) | s[i]; return v; }
/* U512::from_slice + divrem, src/arith.rs:82-138: bit-serial long division of the 512-bit big-endian value by q.
 * Returns 0 when the quotient does not fit below q (the reference's `None`). */
static int u512_divrem_q(const uint8_t s[64], u256 *quot, u256 *rem) {
    u128 n[4];                                  /* n[3] most significant, :92-95 */
    for (int l = 3, i = 0; l >= 0; l--, i += 16) n[l] = be128(s + i);
    u256 q = {{0, 0}}, r = {{0, 0}};
    int q_some = 1;
    for (int i = 511; i >= 0; i--) {
        r.l[1] = (r.l[1] << 1) | (r.l[0] >> 127);   /* mul2; the modulus' two top bits are clear */
        r.l[0] = (r.l[0] << 1) | ((n[i >> 7] >> (i & 127)) & 1);
        if (u256_cmp(&r, &FQ.m) >= 0) {
            sub_noborrow(&r, &FQ.m);
            if (i >= 256) q_some = 0; else q.l[i >> 7] |= (u128)1 << (i & 127);
        }
    }
    *rem = r;
    if (!q_some || u256_cmp(&q, &FQ.m) >= 0) return 0;
    *quot = q;
    return 1;
}
static fq fq_from_u256(u256 a) { u256_mont_mul(&a, &FQ.r2, &FQ.m, FQ.inv); return a; }   /* Fq::new for a < q */

/* G1::from_compressed, src/lib.rs:359-375.  in: n x 33 bytes.  out: n x G1 (z = one), zero-filled on error.
 * status: 0 Ok, 1 CurveError::InvalidEncoding, 2 CurveError::NotMember, 3 CurveError::Field(NotMember).
 * The checks run in the reference's order: field membership, square root, tag, AffineG1::new. */
void orc_g1_from_compressed_many(const uint8_t *in, void *out, uint8_t *status, size_t n) {
    orc_init();
    fq b = fq_from_u64(3);
    for (size_t i = 0; i < n; i++) {
        const uint8_t *s = in + 33 * i;
        g1 res;
        memset(&res, 0, sizeof res);
        uint8_t st = 0;
        u256 xv = {{be128(s + 17), be128(s + 1)}};          /* U256::from_slice: big-endian */
        if (u256_cmp(&xv, &FQ.m) >= 0) st = 3;
        else {
            fq x = fq_from_u256(xv), y;
            fq y2 = fq_add(fq_mul(fq_mul(x, x), x), b);
            if (!fq_sqrt(y2, &y)) st = 2;
            else {
                int odd = (int)(fq_canonical(y).l[0] & 1);
                if (s[0] == 2) { if (odd) y = fq_neg(y); }
                else if (s[0] == 3) { if (!odd) y = fq_neg(y); }
                else st = 1;
                if (!st) {
                    if (!fq_eq(fq_sqr(y), fq_add(fq_mul(fq_sqr(x), x), b))) st = 2;
                    else { res.x = x; res.y = y; res.z = FQ.one; }
                }
            }
        }
        status[i] = st;
        ST(g1, out, i, res);
    }
}
/* G2::from_compressed, src/lib.rs:506-526 (Fq2::from_slice :260-267, to_u512 src/fields/fq2.rs:226-231). */
void orc_g2_from_compressed_many(const uint8_t *in, void *out, uint8_t *status, size_t n, int threads) {
    orc_init();
    threads = clamp_threads(threads);
    u256 rm1 = FR.m;
    rm1.l[0] -= 1;
#pragma omp parallel for num_threads(threads) schedule(dynamic, 1)
    for (size_t i = 0; i < n; i++) {
        const uint8_t *s = in + 65 * i;
        g2 res;
        memset(&res, 0, sizeof res);
        uint8_t st = 0;
        u256 c1, c0;
        if (!u512_divrem_q(s + 1, &c1, &c0)) st = 3;
        else {
            fq2 x = {fq_from_u256(c0), fq_from_u256(c1)}, y;
            fq2 y2 = fq2_add(fq2_mul(fq2_mul(x, x), x), G2_B);
            if (!fq2_sqrt(y2, &y)) st = 2;
            else {
                fq2 yn = fq2_neg(y);
                u256 a1 = fq_canonical(y.c1), a0 = fq_canonical(y.c0), b1 = fq_canonical(yn.c1), b0 = fq_canonical(yn.c0);
                int c = u256_cmp(&a1, &b1);     /* c1 * q + c0 compares lexicographically on (c1, c0) */
                int y_gt = c > 0 || (c == 0 && u256_cmp(&a0, &b0) > 0);
                fq2 e_y = y;
                if (s[0] == 10) e_y = y_gt ? yn : y;
                else if (s[0] == 11) e_y = y_gt ? y : yn;
                else st = 1;
                if (!st) {
                    if (!fq2_eq(fq2_sqr(e_y), fq2_add(fq2_mul(fq2_sqr(x), x), G2_B))) st = 2;
                    else {
                        g2 p;
                        p.x = x; p.y = e_y; p.z = FQ2_ONE;
                        g2 t = g2_add(g2_mul(p, &rm1), p);
                        if (!fq2_is_zero(t.z)) st = 2; else res = p;
                    }
                }
            }
        }
        status[i] = st;
        ST(g2, out, i, res);
    }
}
/* test hook for the divrem restatement: 64 big-endian bytes -> quotient and remainder as little-endian u256 images */
int orc_u512_divrem_q(const uint8_t *be, void *quot, void *rem) {
    orc_init();
    u256 q = {{0, 0}}, r;
    int ok = u512_divrem_q(be, &q, &r);
    memcpy(quot, &q, 32);
    memcpy(rem, &r, 32);
    return ok;
}
void orc_fq_sqrt_many(const void *a, void *out, uint8_t *some, size_t n) {
    orc_init();
    for (size_t i = 0; i < n; i++) { fq r = FQ_ZERO; some[i] = (uint8_t)fq_sqrt(LD(fq, a, i), &r); ST(fq, out, i, r); }
}
void orc_fq2_sqrt_many(const void *a, void *out, uint8_t *some, size_t n) {
    orc_init();
    for (size_t i = 0; i < n; i++) { fq2 r = FQ2_ZERO; some[i] = (uint8_t)fq2_sqrt(LD(fq2, a, i), &r); ST(fq2, out, i, r); }
}
void orc_g1_add_many(const void *a, const void *b, void *out, size_t n) { orc_init(); for (size_t i = 0; i < n; i++) ST(g1, out, i, g1_add(LD(g1, a, i), LD(g1, b, i))); }
void orc_g2_add_many(const void *a, const void *b, void *out, size_t n) { orc_init(); for (size_t i = 0; i < n; i++) ST(g2, out, i, g2_add(LD(g2, a, i), LD(g2, b, i))); }
void orc_g1_double_many(const void *a, void *out, size_t n) { orc_init(); for (size_t i = 0; i < n; i++) ST(g1, out, i, g1_double(LD(g1, a, i))); }
void orc_g2_double_many(const void *a, void *out, size_t n) { orc_init(); for (size_t i = 0; i < n; i++) ST(g2, out, i, g2_double(LD(g2, a, i))); }
/* Group::normalize, lib.rs:391-398 / :542-549: infinity is left untouched */
void orc_g1_normalize_many(const void *p, void *out, size_t n) {
    orc_init();
    for (size_t i = 0; i < n; i++) { g1 v = LD(g1, p, i); g1a a; if (g1_to_affine(v, &a)) { v.x = a.x; v.y = a.y; v.z = FQ.one; } ST(g1, out, i, v); }
}
void orc_g2_normalize_many(const void *p, void *out, size_t n) {
    orc_init();
    for (size_t i = 0; i < n; i++) { g2 v = LD(g2, p, i); g2a a; if (g2_to_affine(v, &a)) { v.x = a.x; v.y = a.y; v.z = FQ2_ONE; } ST(g2, out, i, v); }
}

/* pairing ---------------------------------------------------------------------------------------- */
/* G2Precomp image: q (128 B) followed by 87 x EllCoeffs (192 B each) = 16,832 B.  Returns 0 if q is infinity. */
int orc_g2_precompute(const void *q, void *out) {
    orc_init();
    g2a qa;
    if (!g2_to_affine(LD(g2, q, 0), &qa)) return 0;
    g2precomp pc;
    g2_precompute(qa, &pc);
    memcpy(out, &pc, sizeof pc);
    return 1;
}
size_t orc_g2precomp_size(void) { return sizeof(g2precomp); }

void orc_pairing_many(const void *p, const void *q, void *out, size_t n, int threads) {     /* lib.rs:611-613 */
    orc_init();
    threads = clamp_threads(threads);
#pragma omp parallel for num_threads(threads) schedule(dynamic, 16)
    for (size_t i = 0; i < n; i++) ST(fq12, out, i, pairing(LD(g1, p, i), LD(g2, q, i)));
}
/* groups::pairing_batch, src/groups/mod.rs:904-926, including the doubled to_affine */
static fq12 pairing_batch(const uint8_t *p, const uint8_t *q, size_t n, g1a *pas, g2precomp *pcs) {
    size_t kept = 0;
    for (size_t i = 0; i < n; i++) {
        g1 pp = LD(g1, p, i); g2 qq = LD(g2, q, i);
        g1a pa; g2a qa;
        int exists = g1_to_affine(pp, &pa) & g2_to_affine(qq, &qa);
        if (exists) {
            g1_to_affine(pp, &pas[kept]);
            g2_to_affine(qq, &qa);
            g2_precompute(qa, &pcs[kept]);
            kept++;
        }
    }
    if (kept == 0) return FQ12_ONE;
    fq12 out = FQ12_ONE;
    fq12_final_exponentiation(miller_loop_batch(pcs, pas, kept), &out);
    return out;
}
#include <stdlib.h>
/* n_checks independent pairing_batch calls of pairs_per_check pairs each; is_one[i] = (result == Gt::one()) */
int orc_pairing_product_many(const void *p, const void *q, size_t pairs_per_check, size_t n_checks, void *out,
                             uint8_t *is_one, int threads) {
    orc_init();
    threads = clamp_threads(threads);
    int fail = 0;
#pragma omp parallel num_threads(threads)
    {
        g1a *pas = (g1a *)malloc(sizeof(g1a) * (pairs_per_check ? pairs_per_check : 1));
        g2precomp *pcs = (g2precomp *)malloc(sizeof(g2precomp) * (pairs_per_check ? pairs_per_check : 1));
        if (!pas || !pcs) {
#pragma omp atomic write
            fail = 1;
        } else {
#pragma omp for schedule(dynamic, 4)
            for (size_t i = 0; i < n_checks; i++) {
                fq12 r = pairing_batch((const uint8_t *)p + i * pairs_per_check * sizeof(g1),
                                       (const uint8_t *)q + i * pairs_per_check * sizeof(g2), pairs_per_check, pas, pcs);
                if (out) ST(fq12, out, i, r);
                if (is_one) is_one[i] = (uint8_t)(memcmp(&r, &FQ12_ONE, sizeof r) == 0);
            }
        }
        free(pas); free(pcs);
    }
    return fail ? -1 : 0;
}
/* lib::miller_loop_batch, src/lib.rs:625-633: (G2, G1) order; returns 1 = Ok, 0 = Err(ToAffineConversion) */
int orc_miller_loop_batch(const void *q, const void *p, size_t n, void *out) {
    orc_init();
    g1a *pas = (g1a *)malloc(sizeof(g1a) * (n ? n : 1));
    g2precomp *pcs = (g2precomp *)malloc(sizeof(g2precomp) * (n ? n : 1));
    int ok = pas && pcs;
    for (size_t i = 0; ok && i < n; i++) {
        g2a qa;
        if (!g2_to_affine(LD(g2, q, i), &qa)) { ok = 0; break; }
        g2_precompute(qa, &pcs[i]);
        if (!g1_to_affine(LD(g1, p, i), &pas[i])) { ok = 0; break; }
    }
    if (ok) ST(fq12, out, 0, miller_loop_batch(pcs, pas, n));
    free(pas); free(pcs);
    return ok;
}
/* many independent single-pair Miller loops (pre-final-exp), for checking the fused CUDA Miller kernel */
void orc_miller_loop_many(const void *q, const void *p, void *out, uint8_t *ok, size_t n, int threads) {
    orc_init();
    threads = clamp_threads(threads);
#pragma omp parallel for num_threads(threads) schedule(dynamic, 16)
    for (size_t i = 0; i < n; i++) {
        g1a pa; g2a qa; fq12 r = FQ12_ONE;
        int good = g1_to_affine(LD(g1, p, i), &pa) & g2_to_affine(LD(g2, q, i), &qa);
        if (good) { g2precomp pc; g2_precompute(qa, &pc); r = miller_loop_batch(&pc, &pa, 1); }
        if (ok) ok[i] = (uint8_t)good;
        ST(fq12, out, i, r);
    }
}
/* Montgomery conversion helpers for the harness */
void orc_fq_from_canonical(const void *a, void *out, size_t n) { orc_init(); for (size_t i = 0; i < n; i++) { fq v = LD(fq, a, i); u256_mont_mul(&v, &FQ.r2, &FQ.m, FQ.inv); ST(fq, out, i, v); } }
void orc_fq_to_canonical(const void *a, void *out, size_t n) { orc_init(); u256 one = {{1, 0}}; for (size_t i = 0; i < n; i++) { fq v = LD(fq, a, i); u256_mont_mul(&v, &one, &FQ.m, FQ.inv); ST(fq, out, i, v); } }
void orc_fr_from_canonical(const void *a, void *out, size_t n) { orc_init(); for (size_t i = 0; i < n; i++) { u256 v = LD(u256, a, i); u256_mont_mul(&v, &FR.r2, &FR.m, FR.inv); ST(u256, out, i, v); } }
void orc_fr_to_canonical(const void *a, void *out, size_t n) { orc_init(); u256 one = {{1, 0}}; for (size_t i = 0; i < n; i++) { u256 v = LD(u256, a, i); u256_mont_mul(&v, &one, &FR.m, FR.inv); ST(u256, out, i, v); } }
