// SURVEY 8f rows N2 + N3: AffineG::new validation, Fq/Fq2 square roots and the compressed wire formats
// (G1::from_compressed src/lib.rs:359-375, G2::from_compressed :506-526).  One thread per record.
// Every function is host-compilable (tests/hostsim) and touches only statically named locals: no pointer
// tables, no nullable pointers (DESIGN F7: either would push the whole module to generic LD/ST).
#pragma once
#include "bn_curve.cuh"

namespace bn {

#define BN254_DECODE_OK_ 0
#define BN254_DECODE_INVALID_ENCODING_ 1
#define BN254_DECODE_NOT_MEMBER_ 2
#define BN254_DECODE_FIELD_NOT_MEMBER_ 3

// a >= q on raw little-endian limbs (U256 ordering, src/arith.rs:243-259)
BN_HD bool limbs_ge_q(const uint32_t (&a)[8]) {
    const uint32_t ql[8] = BN_Q_LIMBS;
    uint32_t borrow = 0;
#pragma unroll
    for (int j = 0; j < 8; j++) {
        uint64_t d = (uint64_t)a[j] - ql[j] - borrow;
        borrow = (uint32_t)(d >> 63);
    }
    return borrow == 0;
}
// Fq::new for a value already known to be below q: Montgomery image a * R (src/fields/fp.rs:46-54)
BN_HD void fq_from_canonical(Fq& r, const uint32_t (&a)[8]) {
    Fq t, r2 = BN_C_R2;
#pragma unroll
    for (int j = 0; j < 8; j++) t.v[j] = a[j];
    fq_mul_ni(r, t, r2);
}
// From<Fq> for U256 (src/fields/fp.rs:13-20): Montgomery multiplication by the raw integer 1
BN_HD void fq_to_canonical(uint32_t (&out)[8], const Fq& a) {
    Fq one_raw, t;
    fq_set_zero(one_raw);
    one_raw.v[0] = 1u;
    fq_mul_ni(t, a, one_raw);
#pragma unroll
    for (int j = 0; j < 8; j++) out[j] = t.v[j];
}
// 32 big-endian bytes -> little-endian limbs (U256::from_slice, src/arith.rs:150-165)
BN_HD void limbs_from_be(uint32_t (&out)[8], const uint8_t* s) {
#pragma unroll
    for (int j = 0; j < 8; j++) {
        const uint8_t* p = s + 28 - 4 * j;
        out[j] = ((uint32_t)p[0] << 24) | ((uint32_t)p[1] << 16) | ((uint32_t)p[2] << 8) | (uint32_t)p[3];
    }
}
// (q - 3) / 4 and (q - 1) / 2, the exponents at src/fields/fp.rs:233-241 as integers (q = 3 mod 4)
BN_HD void exp_q_minus3_div4(uint32_t (&e)[8]) {
    uint32_t t[8] = BN_Q_LIMBS;
    t[0] -= 3u;                                   // low limb ...fd47: no borrow
#pragma unroll
    for (int j = 0; j < 8; j++) e[j] = (t[j] >> 2) | (j < 7 ? t[j + 1] << 30 : 0u);
}
BN_HD void exp_q_minus1_div2(uint32_t (&e)[8]) {
    uint32_t t[8] = BN_Q_LIMBS;
    t[0] -= 1u;
#pragma unroll
    for (int j = 0; j < 8; j++) e[j] = (t[j] >> 1) | (j < 7 ? t[j + 1] << 31 : 0u);
}
// FieldElement::pow (src/fields/mod.rs:35-46) for an exponent whose top set bit is `top`: the reference's
// leading squarings of one are skipped, the value is the same.
BN_HD_NOINLINE void fq_pow_limbs(Fq& r, const Fq& a, const uint32_t (&e)[8], int top) {
    Fq acc = a;
    for (int i = top - 1; i >= 0; i--) {
        fq_mul_ni(acc, acc, acc);
        if ((e[i >> 5] >> (i & 31)) & 1) fq_mul_ni(acc, acc, a);
    }
    r = acc;
}
BN_HD_NOINLINE void fq2_pow_limbs(Fq2& r, const Fq2& a, const uint32_t (&e)[8], int top) {
    Fq2 acc = a;
    for (int i = top - 1; i >= 0; i--) {
        fq2_sqr(acc, acc);
        if ((e[i >> 5] >> (i & 31)) & 1) fq2_mul(acc, acc, a);
    }
    r = acc;
}
// Fq::sqrt, src/fields/fp.rs:245-260
BN_HD bool fq_sqrt(Fq& r, const Fq& a) {
    uint32_t e[8];
    exp_q_minus3_div4(e);
    Fq a1, a1a, a0, m1, one = fq_one();
    fq_pow_limbs(a1, a, e, 251);
    fq_mul_ni(a1a, a1, a);
    fq_mul_ni(a0, a1, a1a);
    fq_neg(m1, one);
    if (fq_eq(a0, m1)) return false;
    r = a1a;
    return true;
}
// Fq2::sqrt, src/fields/fq2.rs:208-224.  alpha^q is the Frobenius conjugate, so a0 = alpha^q * alpha is the
// norm c0^2 + c1^2 (imaginary part zero) -- the same element the reference reaches with a 254-bit pow.
BN_HD bool fq2_sqrt(Fq2& r, const Fq2& a) {
    uint32_t e[8];
    exp_q_minus3_div4(e);
    Fq2 a1, a1a, alpha;
    fq2_pow_limbs(a1, a, e, 251);
    fq2_mul(a1a, a1, a);
    fq2_mul(alpha, a1, a1a);
    Fq n, t, m1, one = fq_one();
    fq_mul_ni(n, alpha.c0, alpha.c0);
    fq_mul_ni(t, alpha.c1, alpha.c1);
    fq_add(n, n, t);
    fq_neg(m1, one);
    if (fq_eq(n, m1)) return false;
    if (fq_eq(alpha.c0, m1) && fq_is_zero(alpha.c1)) {          // alpha == -1: i * a1a
        Fq2 i;
        fq_set_zero(i.c0);
        i.c1 = one;
        fq2_mul(r, i, a1a);
    } else {
        Fq2 b, ap1 = alpha;
        fq_add(ap1.c0, alpha.c0, one);
        exp_q_minus1_div2(e);
        fq2_pow_limbs(b, ap1, e, 252);
        fq2_mul(r, b, a1a);
    }
    return true;
}
// Some square root of a in Fq2 (or false), with two Fq exponentiations instead of two Fq2 ones.  Used only by
// G2::from_compressed, whose result does not depend on WHICH root sqrt returns: the tag picks between y and -y by
// comparing them (src/lib.rs:515-523).  Norm method for q = 3 mod 4: n = a0^2 + a1^2 must be a residue in Fq (else
// a is a non-square in Fq2); with s = sqrt(n), exactly one of t = (a0 + s)/2 and t - s is a residue when a1 != 0
// (their product is -(a1/2)^2 and -1 is a non-residue).  r = t^((q-3)/4), x = r t satisfies x^2 = chi(t) t and
// x r = chi(t), so 1/x = chi(t) r: no inversion.  chi(t) = 1: root (x, a1 r / 2); chi(t) = -1: root (-a1 r / 2, x).
BN_HD bool fq2_sqrt_any(Fq2& r, const Fq2& a) {
    uint32_t e[8];
    exp_q_minus3_div4(e);
    Fq n, t, w, s, x, h, one = fq_one();
    if (fq_is_zero(a.c1)) {                         // a in Fq: sqrt(a0) or u * sqrt(-a0)
        fq_pow_limbs(w, a.c0, e, 251);
        fq_mul_ni(x, w, a.c0);
        fq_mul_ni(t, x, x);
        if (fq_is_zero(a.c0)) { r = fq2_zero(); return true; }
        if (fq_eq(t, a.c0)) { r.c0 = x; fq_set_zero(r.c1); } else { fq_set_zero(r.c0); r.c1 = x; }
        return true;
    }
    fq_mul_ni(n, a.c0, a.c0);
    fq_mul_ni(t, a.c1, a.c1);
    fq_add(n, n, t);
    fq_pow_limbs(w, n, e, 251);
    fq_mul_ni(s, w, n);                             // candidate sqrt(n)
    fq_mul_ni(t, s, s);
    if (!fq_eq(t, n)) return false;                 // the norm is a non-residue: no square root in Fq2
    fq_add(t, a.c0, s);
    fq_half(t, t);                                  // t = (a0 + s) / 2, non-zero because a1 != 0
    fq_pow_limbs(w, t, e, 251);
    fq_mul_ni(x, w, t);                             // x^2 = +-t
    fq_mul_ni(h, a.c1, w);
    fq_half(h, h);                                  // a1 r / 2
    fq_mul_ni(n, x, x);
    if (fq_eq(n, t)) { r.c0 = x; r.c1 = h; }
    else { fq_neg(r.c0, h); r.c1 = x; }
    (void)one;
    return true;
}
// U512::from_slice + divrem(q), src/arith.rs:82-138: 64 big-endian bytes n -> (n / q, n % q).  false when
// the quotient is not below q (the reference's None).  Bit-serial like the reference; the word index of the
// dividend is static (outer loop unrolled), only the bit loop runs.
BN_HD_NOINLINE bool u512_divrem_q(uint32_t (&quot)[8], uint32_t (&rem)[8], const uint8_t* s) {
    const uint32_t ql[8] = BN_Q_LIMBS;
    uint32_t r[8];
#pragma unroll
    for (int j = 0; j < 8; j++) { r[j] = 0; quot[j] = 0; }
    bool fits = true;
#pragma unroll
    for (int w = 15; w >= 0; w--) {
        const uint8_t* p = s + 60 - 4 * w;
        uint32_t word = ((uint32_t)p[0] << 24) | ((uint32_t)p[1] << 16) | ((uint32_t)p[2] << 8) | (uint32_t)p[3];
        uint32_t qw = 0;
#pragma unroll 1
        for (int b = 0; b < 32; b++) {
#pragma unroll
            for (int j = 7; j > 0; j--) r[j] = (r[j] << 1) | (r[j - 1] >> 31);
            r[0] = (r[0] << 1) | (word >> 31);
            word <<= 1;
            uint32_t t[8], borrow = 0;
#pragma unroll
            for (int j = 0; j < 8; j++) {
                uint64_t d = (uint64_t)r[j] - ql[j] - borrow;
                t[j] = (uint32_t)d;
                borrow = (uint32_t)(d >> 63);
            }
            uint32_t ge = borrow ^ 1u;
            if (ge) {
#pragma unroll
                for (int j = 0; j < 8; j++) r[j] = t[j];
            }
            qw = (qw << 1) | ge;
        }
        if (w >= 8) { if (qw) fits = false; } else quot[w] = qw;
    }
#pragma unroll
    for (int j = 0; j < 8; j++) rem[j] = r[j];
    return fits && !limbs_ge_q(quot);
}

// AffineG::new, src/groups/mod.rs:94-113
BN_HD bool g1_affine_on_curve(const AffG1& a) {
    Fq lhs, rhs, b = BN_C_G1_B;
    fq_mul_ni(lhs, a.y, a.y);
    fq_mul_ni(rhs, a.x, a.x);
    fq_mul_ni(rhs, rhs, a.x);
    fq_add(rhs, rhs, b);
    return fq_eq(lhs, rhs);
}
BN_HD bool g2_affine_on_curve(const AffG2& a) {
    Fq2 lhs, rhs, b = BN_C_G2_B;
    fq2_sqr(lhs, a.y);
    fq2_sqr(rhs, a.x);
    fq2_mul(rhs, rhs, a.x);
    fq2_add(rhs, rhs, b);
    return fq2_eq(lhs, rhs);
}
// the order check at :97-107: p * (-Fr::one()) + p == 0, i.e. [r - 1]p + p with the reference's own ladder
BN_HD_NOINLINE bool g2_in_subgroup_ref(const AffG2& a) {
    uint32_t k[8] = BN_R_LIMBS;
    k[0] -= 1u;                                   // r is odd: no borrow
    G2 p, t;
    p.x = a.x; p.y = a.y; p.z = fq2_one();
    jac_mul(t, p, k);
    jac_add(t, t, p);
    return fq2_is_zero(t.z);
}
// The same predicate without the 254-bit ladder.  psi = twist^-1 . Frobenius . twist (AffineG2::mul_by_q, :694-699,
// here on Jacobian triples: conjugate X, Y, Z and scale X, Y by the two twist constants) satisfies
// psi^2 - t psi + q = 0 on all of E'(Fq2) and acts on the order-r subgroup as multiplication by q = 6u^2 mod r.
// For BN curves  [u+1]P + psi([u]P) + psi^2([u]P) = psi^3([2u]P)  holds on that subgroup; conversely a point of
// E'(Fq2) satisfying it is killed by the resultant of f(X) = -2uX^3 + uX^2 + uX + u + 1 and X^2 - tX + q, an
// integer that r divides and that is coprime to the cofactor 2q - r (checked numerically in
// tests/test_wire_oracle.py::test_fast_subgroup_test_is_exact), so its order divides r: the two predicates
// agree on every on-curve input, including the small-order points of the cofactor.  One 63-bit ladder
// instead of a 254-bit one.
BN_HD void g2_psi(G2& r, const G2& p) {
    const Fq2 cx = BN_C_TWIST_MUL_BY_Q_X, cy = BN_C_TWIST_MUL_BY_Q_Y;
    Fq2 t;
    fq2_conj(t, p.x); fq2_mul(r.x, t, cx);
    fq2_conj(t, p.y); fq2_mul(r.y, t, cy);
    fq2_conj(r.z, p.z);
}
BN_HD_NOINLINE bool g2_in_subgroup_psi(const AffG2& a) {
    uint32_t k[8] = {BN_U_LO, BN_U_HI, 0u, 0u, 0u, 0u, 0u, 0u};
    G2 p, up, b, c, d;
    p.x = a.x; p.y = a.y; p.z = fq2_one();
    jac_mul(up, p, k);                            // [u]P
    g2_psi(b, up);                                // psi([u]P)
    g2_psi(c, b);                                 // psi^2([u]P)
    g2_psi(d, c);                                 // psi^3([u]P)
    jac_add(up, up, p);                           // [u+1]P
    jac_add(up, up, b);
    jac_add(up, up, c);                           // left-hand side
    jac_double(d, d);                             // psi^3([2u]P)
    fq2_neg(up.y, up.y);
    jac_add(d, d, up);
    return fq2_is_zero(d.z);
}
BN_HD uint8_t g1_affine_status(const AffG1& a) { return g1_affine_on_curve(a) ? 0 : 1; }
BN_HD uint8_t g2_affine_status(const AffG2& a, bool ref_ladder) {
    if (!g2_affine_on_curve(a)) return 1;
    bool in = ref_ladder ? g2_in_subgroup_ref(a) : g2_in_subgroup_psi(a);
    return in ? 0 : 2;
}

// G1::from_compressed, src/lib.rs:359-375: checks in the reference's order (field, root, tag, AffineG1::new)
BN_HD uint8_t g1_decode(G1& out, const uint8_t* s) {
    uint32_t xv[8];
    limbs_from_be(xv, s + 1);
    if (limbs_ge_q(xv)) return BN254_DECODE_FIELD_NOT_MEMBER_;
    AffG1 a;
    fq_from_canonical(a.x, xv);
    Fq y2, b = BN_C_G1_B;
    fq_mul_ni(y2, a.x, a.x);
    fq_mul_ni(y2, y2, a.x);
    fq_add(y2, y2, b);
    if (!fq_sqrt(a.y, y2)) return BN254_DECODE_NOT_MEMBER_;
    uint32_t yc[8];
    fq_to_canonical(yc, a.y);
    bool odd = yc[0] & 1u;
    uint8_t tag = s[0];
    if (tag != 2 && tag != 3) return BN254_DECODE_INVALID_ENCODING_;
    if ((tag == 2) == odd) fq_neg(a.y, a.y);
    if (!g1_affine_on_curve(a)) return BN254_DECODE_NOT_MEMBER_;
    out.x = a.x; out.y = a.y; out.z = fq_one();
    return BN254_DECODE_OK_;
}
// a > b as the integers c1 * q + c0 (Fq2::to_u512, src/fields/fq2.rs:226-231): lexicographic on canonical (c1, c0)
BN_HD int limbs_cmp(const uint32_t (&a)[8], const uint32_t (&b)[8]) {
    for (int j = 7; j >= 0; j--) {
        if (a[j] != b[j]) return a[j] > b[j] ? 1 : -1;
    }
    return 0;
}
BN_HD bool fq2_u512_gt(const Fq2& a, const Fq2& b) {
    uint32_t x[8], y[8];
    fq_to_canonical(x, a.c1);
    fq_to_canonical(y, b.c1);
    int c = limbs_cmp(x, y);
    if (c != 0) return c > 0;
    fq_to_canonical(x, a.c0);
    fq_to_canonical(y, b.c0);
    return limbs_cmp(x, y) > 0;
}
// G2::from_compressed, src/lib.rs:506-526
BN_HD uint8_t g2_decode(G2& out, const uint8_t* s, bool ref_ladder) {
    uint32_t c1[8], c0[8];
    if (!u512_divrem_q(c1, c0, s + 1)) return BN254_DECODE_FIELD_NOT_MEMBER_;
    AffG2 a;
    fq_from_canonical(a.x.c0, c0);
    fq_from_canonical(a.x.c1, c1);
    Fq2 y2, y, yn, b = BN_C_G2_B;
    fq2_sqr(y2, a.x);
    fq2_mul(y2, y2, a.x);
    fq2_add(y2, y2, b);
    if (!fq2_sqrt_any(y, y2)) return BN254_DECODE_NOT_MEMBER_;
    fq2_neg(yn, y);
    bool y_gt = fq2_u512_gt(y, yn);
    uint8_t tag = s[0];
    if (tag != 10 && tag != 11) return BN254_DECODE_INVALID_ENCODING_;
    a.y = ((tag == 10) == y_gt) ? yn : y;
    uint8_t st = g2_affine_status(a, ref_ladder);
    if (st != 0) return BN254_DECODE_NOT_MEMBER_;
    out.x = a.x; out.y = a.y; out.z = fq2_one();
    return BN254_DECODE_OK_;
}

}  // namespace bn
