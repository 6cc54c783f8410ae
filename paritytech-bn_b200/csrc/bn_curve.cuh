// G1 / G2 Jacobian group law, scalar multiplication, G2 line steps and the optimal-ate Miller loop.
// Follows src/groups/mod.rs of the reference (file:line cited per function); field work is the lean
// evaluation in bn_tower.cuh.
#pragma once
#include "bn_tower.cuh"

namespace bn {

// ---------------------------------------------------------------------------------------------
// field-op adaptors so the group law is written once for Fq (G1) and Fq2 (G2)
// ---------------------------------------------------------------------------------------------
struct FqOps {
    typedef Fq T;
    static BN_HD void add(T& r, const T& a, const T& b) { fq_add(r, a, b); }
    static BN_HD void sub(T& r, const T& a, const T& b) { fq_sub(r, a, b); }
    static BN_HD void neg(T& r, const T& a) { fq_neg(r, a); }
    static BN_HD void mul(T& r, const T& a, const T& b) { fq_mul_ni(r, a, b); }
    static BN_HD void sqr(T& r, const T& a) { fq_mul_ni(r, a, a); }
    static BN_HD bool inv(T& r, const T& a) { return fq_inv_bgcd(r, a); }
    static BN_HD bool is_zero(const T& a) { return fq_is_zero(a); }
    static BN_HD bool eq(const T& a, const T& b) { return fq_eq(a, b); }
    static BN_HD T zero() { return fq_zero(); }
    static BN_HD T one() { return fq_one(); }
};
struct Fq2Ops {
    typedef Fq2 T;
    static BN_HD void add(T& r, const T& a, const T& b) { fq2_add(r, a, b); }
    static BN_HD void sub(T& r, const T& a, const T& b) { fq2_sub(r, a, b); }
    static BN_HD void neg(T& r, const T& a) { fq2_neg(r, a); }
    static BN_HD void mul(T& r, const T& a, const T& b) { fq2_mul(r, a, b); }
    static BN_HD void sqr(T& r, const T& a) { fq2_sqr(r, a); }
    static BN_HD bool inv(T& r, const T& a) { return fq2_inv(r, a); }
    static BN_HD bool is_zero(const T& a) { return fq2_is_zero(a); }
    static BN_HD bool eq(const T& a, const T& b) { return fq2_eq(a, b); }
    static BN_HD T zero() { return fq2_zero(); }
    static BN_HD T one() { return fq2_one(); }
};

template <class F> struct Jac { typename F::T x, y, z; };   // G<P>, src/groups/mod.rs:45-50 (repr(C))
template <class F> struct Aff { typename F::T x, y; };      // AffineG<P>, :82-86
typedef Jac<FqOps> G1;
typedef Jac<Fq2Ops> G2;
typedef Aff<FqOps> AffG1;
typedef Aff<Fq2Ops> AffG2;

template <class F> BN_HD void jac_set_zero(Jac<F>& p) {     // (0, 1, 0), :230-236
    p.x = F::zero(); p.y = F::one(); p.z = F::zero();
}
// dbl-2009-l, src/groups/mod.rs:250-269
template <class F> BN_HD_NOINLINE void jac_double(Jac<F>& r, const Jac<F>& p) {
    typename F::T a, b, c, d, e, f, x3, t, y1z1;
    F::sqr(a, p.x);
    F::sqr(b, p.y);
    F::sqr(c, b);
    F::add(t, p.x, b); F::sqr(d, t); F::sub(d, d, a); F::sub(d, d, c);
    F::add(d, d, d);
    F::add(e, a, a); F::add(e, e, a);
    F::sqr(f, e);
    F::add(t, d, d); F::sub(x3, f, t);
    F::add(c, c, c); F::add(c, c, c); F::add(c, c, c);      // 8c
    F::mul(y1z1, p.y, p.z);
    F::sub(t, d, x3); F::mul(t, e, t); F::sub(r.y, t, c);
    r.x = x3;
    F::add(r.z, y1z1, y1z1);
}
// add-2007-bl with the reference's zero / equal-point branches, src/groups/mod.rs:294-334
template <class F> BN_HD_NOINLINE void jac_add(Jac<F>& r, const Jac<F>& p, const Jac<F>& o) {
    if (F::is_zero(p.z)) { r = o; return; }
    if (F::is_zero(o.z)) { r = p; return; }
    typename F::T z1s, z2s, u1, u2, z1c, z2c, s1, s2, h, sm, i, j, rr, v, s1j, x3, t, zz;
    F::sqr(z1s, p.z);
    F::sqr(z2s, o.z);
    F::mul(u1, p.x, z2s);
    F::mul(u2, o.x, z1s);
    F::mul(z1c, p.z, z1s);
    F::mul(z2c, o.z, z2s);
    F::mul(s1, p.y, z2c);
    F::mul(s2, o.y, z1c);
    if (F::eq(u1, u2) && F::eq(s1, s2)) { jac_double(r, p); return; }
    F::sub(h, u2, u1);
    F::sub(sm, s2, s1);
    F::add(t, h, h); F::sqr(i, t);
    F::mul(j, h, i);
    F::add(rr, sm, sm);
    F::mul(v, u1, i);
    F::mul(s1j, s1, j);
    F::sqr(x3, rr); F::sub(x3, x3, j); F::add(t, v, v); F::sub(x3, x3, t);
    F::add(zz, p.z, o.z); F::sqr(zz, zz); F::sub(zz, zz, z1s); F::sub(zz, zz, z2s);
    F::sub(t, v, x3); F::mul(t, rr, t); F::add(s1j, s1j, s1j); F::sub(r.y, t, s1j);
    r.x = x3;
    F::mul(r.z, zz, h);
}
// Neg, src/groups/mod.rs:336-350: (x, -y, z); the point at infinity is returned unchanged
template <class F> BN_HD void jac_neg(Jac<F>& r, const Jac<F>& p) {
    r = p;
    if (!F::is_zero(p.z)) F::neg(r.y, p.y);
}
// PartialEq, src/groups/mod.rs:169-195: projective equality x1 z2^2 == x2 z1^2 and y1 z2^3 == y2 z1^3; infinity only equals infinity
template <class F> BN_HD bool jac_eq(const Jac<F>& p, const Jac<F>& o) {
    if (F::is_zero(p.z)) return F::is_zero(o.z);
    if (F::is_zero(o.z)) return false;
    typename F::T z1s, z2s, a, b;
    F::sqr(z1s, p.z);
    F::sqr(z2s, o.z);
    F::mul(a, p.x, z2s);
    F::mul(b, o.x, z1s);
    if (!F::eq(a, b)) return false;
    F::mul(z1s, p.z, z1s);
    F::mul(z2s, o.z, z2s);
    F::mul(a, p.y, z2s);
    F::mul(b, o.y, z1s);
    return F::eq(a, b);
}
// G * Fr, src/groups/mod.rs:272-292: MSB-first double-and-add over the canonical (de-Montgomeryed)
// scalar bits, exactly the reference's sequence, so even the raw Jacobian triple matches.
template <class F> BN_HD void jac_mul(Jac<F>& r, const Jac<F>& p, const uint32_t (&k)[8]) {
    Jac<F> res;
    jac_set_zero(res);
    bool found_one = false;
    for (int i = 255; i >= 0; i--) {
        if (found_one) jac_double(res, res);
        if ((k[i >> 5] >> (i & 31)) & 1) {
            found_one = true;
            jac_add(res, res, p);
        }
    }
    r = res;
}
// to_affine, src/groups/mod.rs:198-217: false at infinity; copy if z is exactly one
template <class F> BN_HD bool jac_to_affine(Aff<F>& r, const Jac<F>& p) {
    if (F::is_zero(p.z)) return false;
    if (F::eq(p.z, F::one())) { r.x = p.x; r.y = p.y; return true; }
    typename F::T zi, zi2, zi3;
    F::inv(zi, p.z);
    F::sqr(zi2, zi);
    F::mul(r.x, p.x, zi2);
    F::mul(zi3, zi2, zi);
    F::mul(r.y, p.y, zi3);
    return true;
}
// Group::normalize, src/lib.rs:391-398 / :542-549 (infinity untouched)
template <class F> BN_HD void jac_normalize(Jac<F>& r, const Jac<F>& p) {
    Aff<F> a;
    if (jac_to_affine(a, p)) { r.x = a.x; r.y = a.y; r.z = F::one(); } else { r = p; }
}

// Signed 4-bit window recoding of a canonical scalar k < 2^255: k = sum_i d_i 16^i with d_i in {-7..8}.
// mag holds |d_i| as nibble i (0..8), sgn bit i is set where d_i < 0.  (k < r < 2^254, so the top digit is at most 4.)
BN_HD void wnaf4_fixed_recode(uint32_t (&mag)[8], uint32_t (&sgn)[2], const uint32_t (&k)[8]) {
    uint32_t carry = 0;
    sgn[0] = sgn[1] = 0;
    for (int w = 0; w < 8; w++) {
        uint32_t m = 0;
        for (int j = 0; j < 8; j++) {
            uint32_t d = ((k[w] >> (4 * j)) & 15u) + carry;     // 0..16
            carry = d > 8u ? 1u : 0u;
            uint32_t neg = carry;
            uint32_t a = neg ? 16u - d : d;                      // 0..8
            m |= a << (4 * j);
            int i = w * 8 + j;
            sgn[i >> 5] |= neg << (i & 31);
        }
        mag[w] = m;
    }
}
// k * P for callers that normalise the result: fixed signed 4-bit window over the canonical scalar -- a table of
// 1P..8P (4 doublings + 3 additions), then 63 x (4 doublings + one table addition with a conditional negation).
// Every lane of a warp runs the same sequence (the bit-serial ladder of jac_mul executes the addition for almost
// every bit once 32 lanes are combined).  The Jacobian representative differs from jac_mul's; the group element
// is the same (the a = 0 formulas never use b, so this holds for any input the reference computes on), hence the
// normalised bytes agree.  Returns with z = 0 when the result is the point at infinity.
template <class F> BN_HD void jac_mul_window(Jac<F>& r, const Jac<F>& p, const uint32_t (&k)[8]) {
    Jac<F> tab[8];
    tab[0] = p;
    jac_double(tab[1], p);
    jac_add(tab[2], tab[1], p);
    jac_double(tab[3], tab[1]);
    jac_add(tab[4], tab[3], p);
    jac_double(tab[5], tab[2]);
    jac_add(tab[6], tab[5], p);
    jac_double(tab[7], tab[3]);
    uint32_t mag[8], sgn[2];
    wnaf4_fixed_recode(mag, sgn, k);
    Jac<F> acc;
    jac_set_zero(acc);
    for (int i = 63; i >= 0; i--) {
        if (i != 63) { jac_double(acc, acc); jac_double(acc, acc); jac_double(acc, acc); jac_double(acc, acc); }
        uint32_t a = (mag[i >> 3] >> (4 * (i & 7))) & 15u;
        if (a != 0) {
            Jac<F> t = tab[a - 1];
            if ((sgn[i >> 5] >> (i & 31)) & 1u) F::neg(t.y, t.y);
            jac_add(acc, acc, t);
        }
    }
    r = acc;
}
// (p * k).normalize() with the reference's bytes in every case: a result at infinity (k = 0, p at infinity, or a
// point whose multiple vanishes) is recomputed by the reference's own ladder, whose (x, y, 0) it leaves untouched.
template <class F> BN_HD void jac_mul_normalized(Jac<F>& r, const Jac<F>& p, const uint32_t (&k)[8]) {
    Jac<F> t;
    jac_mul_window(t, p, k);
    if (F::is_zero(t.z)) { jac_mul(r, p, k); return; }
    Aff<F> a;
    jac_to_affine(a, t);
    r.x = a.x; r.y = a.y; r.z = F::one();
}

// Fr: canonical scalar bits from the Montgomery image, U256::from(Fr), src/fields/fp.rs:13-20.
// One Montgomery reduction mod r of the 256-bit value (k*R) -> k.
BN_HD void fr_to_canonical(uint32_t (&out)[8], const uint32_t (&m)[8]) {
    const uint32_t rl[8] = BN_R_LIMBS;
    uint32_t t[17];
    for (int i = 0; i < 8; i++) { t[i] = m[i]; t[8 + i] = 0; }
    t[16] = 0;
    for (int i = 0; i < 8; i++) {
        uint32_t kk = t[i] * BN_RINV32;
        uint64_t c = 0;
        for (int j = 0; j < 8; j++) {
            uint64_t s = (uint64_t)kk * rl[j] + t[i + j] + c;
            t[i + j] = (uint32_t)s;
            c = s >> 32;
        }
        for (int j = i + 8; j < 17; j++) {
            uint64_t s = (uint64_t)t[j] + c;
            t[j] = (uint32_t)s;
            c = s >> 32;
        }
    }
    // result t[8..15] < 2r; one conditional subtraction
    bool ge = true;
    for (int i = 7; i >= 0; i--) {
        if (t[8 + i] != rl[i]) { ge = t[8 + i] > rl[i]; break; }
    }
    int64_t br = 0;
    for (int i = 0; i < 8; i++) {
        int64_t d = (int64_t)t[8 + i] - (ge ? rl[i] : 0u) + br;
        out[i] = (uint32_t)d;
        br = d >> 32;
    }
}

// ---------------------------------------------------------------------------------------------
// G2 line steps (homogeneous projective, "flipped" libff convention) and the Miller loop
// ---------------------------------------------------------------------------------------------
struct EllCoeffs { Fq2 ell_0, ell_vw, ell_vv; };   // field order of src/groups/mod.rs:566-571

// a/2 on the Montgomery image: (a + (a odd ? q : 0)) >> 1.  Same value as a * two_inv (:520-528).
BN_HD void fq_half(Fq& rr, const Fq& aa) {
    Fq a = fq_ld(aa), r;
    uint32_t mask = 0u - (a.v[0] & 1u);
    uint32_t t[9];
    uint64_t c = 0;
    for (int i = 0; i < 8; i++) {
        uint64_t s = (uint64_t)a.v[i] + (fq_modulus_limb(i) & mask) + c;
        t[i] = (uint32_t)s;
        c = s >> 32;
    }
    t[8] = (uint32_t)c;
    for (int i = 0; i < 8; i++) r.v[i] = (t[i] >> 1) | (t[i + 1] << 31);
    fq_st(rr, r);
}
BN_HD void fq2_half(Fq2& r, const Fq2& a) { fq_half(r.c0, a.c0); fq_half(r.c1, a.c1); }

// doubling_step_for_flipped_miller_loop, src/groups/mod.rs:754-776
BN_HD_NOINLINE void g2_doubling_step(G2& r, EllCoeffs& out) {
    const Fq2 b3 = BN_C_G2_B3;   // 3*b' : e = b' * (3 Z^2)
    Fq2 a, b, c, e, f, g, h, i, j, esq, t;
    fq2_mul(a, r.x, r.y); fq2_half(a, a);
    fq2_sqr(b, r.y);
    fq2_sqr(c, r.z);
    fq2_mul(e, b3, c);
    fq2_dbl(f, e); fq2_add(f, f, e);
    fq2_add(g, b, f); fq2_half(g, g);
    fq2_add(t, r.y, r.z); fq2_sqr(h, t); fq2_add(t, b, c); fq2_sub(h, h, t);
    fq2_sub(i, e, b);
    fq2_sqr(j, r.x);
    fq2_sqr(esq, e);
    fq2_sub(t, b, f); fq2_mul(r.x, a, t);
    fq2_sqr(g, g); fq2_dbl(t, esq); fq2_add(t, t, esq); fq2_sub(r.y, g, t);
    fq2_mul(r.z, b, h);
    fq2_mul_by_xi(out.ell_0, i);
    fq2_neg(out.ell_vw, h);
    fq2_dbl(t, j); fq2_add(out.ell_vv, t, j);
}
// mixed_addition_step_for_flipped_miller_loop, src/groups/mod.rs:731-752
BN_HD_NOINLINE void g2_mixed_addition_step(G2& r, const AffG2& base, EllCoeffs& out) {
    Fq2 d, e, f, g, h, i, j, t, u;
    fq2_mul(t, r.z, base.x); fq2_sub(d, r.x, t);
    fq2_mul(t, r.z, base.y); fq2_sub(e, r.y, t);
    fq2_sqr(f, d);
    fq2_sqr(g, e);
    fq2_mul(h, d, f);
    fq2_mul(i, r.x, f);
    fq2_mul(j, r.z, g); fq2_add(j, j, h); fq2_dbl(t, i); fq2_sub(j, j, t);
    fq2_mul(r.x, d, j);
    fq2_sub(t, i, j); fq2_mul(t, e, t); fq2_mul(u, h, r.y); fq2_sub(r.y, t, u);
    fq2_mul(r.z, r.z, h);
    fq2_mul(t, e, base.x); fq2_mul(u, d, base.y); fq2_sub(t, t, u);
    fq2_mul_by_xi(out.ell_0, t);
    fq2_neg(out.ell_vv, e);
    cpv(out.ell_vw, d);
}
// AffineG2::mul_by_q, src/groups/mod.rs:694-699
BN_HD void g2_mul_by_q(AffG2& r, const AffG2& q) {
    const Fq2 cx = BN_C_TWIST_MUL_BY_Q_X, cy = BN_C_TWIST_MUL_BY_Q_Y;
    Fq2 t;
    fq2_conj(t, q.x); fq2_mul(r.x, cx, t);
    fq2_conj(t, q.y); fq2_mul(r.y, cy, t);
}

// ATE_LOOP_COUNT_NAF (src/groups/mod.rs:9-14) packed 2 bits per digit, digit k at bits [2k, 2k+1]
// of word k/16: 0, 1, or 3 (= -1).
#define BN_NAF_W0 0x13033011u   /* digits  0..15 */
#define BN_NAF_W1 0x130400c3u   /* digits 16..31 */
#define BN_NAF_W2 0x3301300cu   /* digits 32..47 */
#define BN_NAF_W3 0x01130c04u   /* digits 48..63 */
BN_HD uint32_t naf_digit(int k) {
    uint32_t w = (k < 16) ? BN_NAF_W0 : (k < 32) ? BN_NAF_W1 : (k < 48) ? BN_NAF_W2 : BN_NAF_W3;
    return (w >> (2 * (k & 15))) & 3u;
}
#define BN_N_COEFFS 87
// bit c set <=> coefficient triple c of G2Precomp.coeffs comes from a doubling step (64 of the 87; the Miller
// loop squares f before each of them).  Derived from the NAF table above; checked by
// tests/test_oracle_golden.py::test_doubling_mask_matches_naf.
#define BN_DBL_MASK_LO 0xb7b7ddbbf76dededull
#define BN_DBL_MASK_HI 0x00000000001db77bull
BN_HD bool coeff_is_doubling(int c) { return (((c < 64) ? (BN_DBL_MASK_LO >> c) : (BN_DBL_MASK_HI >> (c - 64))) & 1ull) != 0; }

// one line evaluated at P and multiplied into f: the call shape at src/groups/mod.rs:589
BN_HD void line_mul(Fq12& f, const EllCoeffs& c, const AffG1& p) {
    Fq2 vw, vv;
    fq2_scale(vw, c.ell_vw, p.y);
    fq2_scale(vv, c.ell_vv, p.x);
    fq12_mul_by_024(f, c.ell_0, vw, vv);
}

// AffineG2::precompute, src/groups/mod.rs:701-728: all 87 coefficient triples, via a sink functor
// so the same walk serves "materialise G2Precomp" and "consume on the fly in the Miller loop".
template <class Sink> BN_HD void g2_walk_lines(const AffG2& q, Sink& sink) {
    G2 r;
    r.x = q.x; r.y = q.y; r.z = fq2_one();
    AffG2 qn;
    qn.x = q.x; fq2_neg(qn.y, q.y);
    EllCoeffs c;
    for (int k = 0; k < 64; k++) {
        g2_doubling_step(r, c);
        sink.dbl(c, k);
        uint32_t d = naf_digit(k);
        if (d != 0) {
            g2_mixed_addition_step(r, d == 1 ? q : qn, c);
            sink.add(c);
        }
    }
    AffG2 q1, q2;
    g2_mul_by_q(q1, q);
    g2_mul_by_q(q2, q1);
    fq2_neg(q2.y, q2.y);
    g2_mixed_addition_step(r, q1, c);
    sink.add(c);
    g2_mixed_addition_step(r, q2, c);
    sink.add(c);
}

struct MillerSink {   // G2Precomp::miller_loop (src/groups/mod.rs:580-606) fused with precompute
    Fq12 f;
    AffG1 p;
    BN_HD void dbl(const EllCoeffs& c, int k) {
        if (k != 0) fq12_sqr(f, f);   // the reference squares 1 in the first iteration: same value
        line_mul(f, c, p);
    }
    BN_HD void add(const EllCoeffs& c) { line_mul(f, c, p); }
};
struct StoreSink {    // materialise the coefficient table (G2Precomp.coeffs)
    EllCoeffs* out;
    int n;
    BN_HD void dbl(const EllCoeffs& c, int) { out[n++] = c; }
    BN_HD void add(const EllCoeffs& c) { out[n++] = c; }
};

// Miller loop of one pair, both points affine; pre-final-exponentiation value, bit-identical to
// q.precompute().miller_loop(&p) because the reference's line formulas are used unchanged.
BN_HD void miller_loop_fused(Fq12& f, const AffG1& p, const AffG2& q) {
    MillerSink s;
    s.f = fq12_one();
    s.p = p;
    g2_walk_lines(q, s);
    f = s.f;
}

// groups::pairing, src/groups/mod.rs:894-902 (infinity -> one)
BN_HD void pairing(Fq12& out, const G1& p, const G2& q) {
    AffG1 pa;
    AffG2 qa;
    if (!jac_to_affine(pa, p) || !jac_to_affine(qa, q)) { out = fq12_one(); return; }
    Fq12 f;
    miller_loop_fused(f, pa, qa);
    fq12_final_exponentiation(out, f);
}

}  // namespace bn
