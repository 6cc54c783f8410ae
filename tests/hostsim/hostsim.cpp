// TEST-ONLY host build of the device headers (bn_fq/bn_tower/bn_curve/bn_wire .cuh) with the portable
// arithmetic back end.  Lets the CPU test-suite check every formula, constant and control path
// of the CUDA code against the oracle without a GPU.  Never linked into the shipped library.
#include <stddef.h>
#include <string.h>
#define BN_HOST_SIM 1
#include "../../paritytech-bn_b200/csrc/bn_curve.cuh"
#include "../../paritytech-bn_b200/csrc/bn_wire.cuh"
using namespace bn;

template <class T> static T ld(const void* p, size_t i) { T v; memcpy(&v, (const char*)p + i * sizeof(T), sizeof(T)); return v; }
template <class T> static void st(void* p, size_t i, const T& v) { memcpy((char*)p + i * sizeof(T), &v, sizeof(T)); }

#define BIN(name, T, fn) extern "C" void hs_##name(const void* a, const void* b, void* o, size_t n) { \
    for (size_t i = 0; i < n; i++) { T x = ld<T>(a, i), y = ld<T>(b, i), r; fn(r, x, y); st(o, i, r); } }
#define UN(name, T, fn) extern "C" void hs_##name(const void* a, void* o, size_t n) { \
    for (size_t i = 0; i < n; i++) { T x = ld<T>(a, i), r; fn(r, x); st(o, i, r); } }
// in-place variants exercise the alias-safety the kernels rely on
#define BIN_ALIAS(name, T, fn) extern "C" void hs_##name##_alias(const void* a, const void* b, void* o, size_t n) { \
    for (size_t i = 0; i < n; i++) { T x = ld<T>(a, i), y = ld<T>(b, i); fn(x, x, y); st(o, i, x); } }

BIN(fq_mul, Fq, fq_mul) BIN(fq_add, Fq, fq_add) BIN(fq_sub, Fq, fq_sub)
BIN(fq2_mul, Fq2, fq2_mul) UN(fq2_sqr, Fq2, fq2_sqr) UN(fq2_mul_by_xi, Fq2, fq2_mul_by_xi) UN(fq2_half, Fq2, fq2_half)
BIN(fq6_mul, Fq6, fq6_mul) UN(fq6_sqr, Fq6, fq6_sqr)
BIN(fq12_mul, Fq12, fq12_mul) UN(fq12_sqr, Fq12, fq12_sqr)
BIN_ALIAS(fq2_mul, Fq2, fq2_mul) BIN_ALIAS(fq6_mul, Fq6, fq6_mul) BIN_ALIAS(fq12_mul, Fq12, fq12_mul)
UN(fq12_cyclotomic_sqr, Fq12, fq12_cyclotomic_sqr) UN(fq12_exp_by_neg_z, Fq12, fq12_exp_by_neg_z)
UN(fq12_final_exp_last_chunk, Fq12, fq12_final_exp_last_chunk)
BIN(g1_add, G1, jac_add) BIN(g2_add, G2, jac_add) UN(g1_double, G1, jac_double) UN(g2_double, G2, jac_double)
UN(g1_normalize, G1, jac_normalize) UN(g2_normalize, G2, jac_normalize)

// lazy sums of the quad-lane executor: out = sum_k +-terms[k] (signs bit k) [+ 9 * sum_k +-xterms[k]] mod q
extern "C" void hs_acc9_sum(const void* terms, unsigned nterms, unsigned signs, const void* xterms, unsigned nx, unsigned xsigns, void* o) {
    Acc9 a, x;
    acc9_zero(a); acc9_zero(x);
    for (unsigned k = 0; k < nterms; k++) acc9_add(a, ld<Fq>(terms, k), (signs >> k) & 1u ? 0xffffffffu : 0u);
    for (unsigned k = 0; k < nx; k++) acc9_add(x, ld<Fq>(xterms, k), (xsigns >> k) & 1u ? 0xffffffffu : 0u);
    Fq r;
    if (nx) { acc9_add9(a, x); acc9_reduce<true>(r, a); } else acc9_reduce<false>(r, a);
    st(o, 0, r);
}

// one component of the cyclotomic squaring's out_c (bn_slots.cuh s_cyc_post): 3 (s +- mx - 10 m) - 2 z
extern "C" void hs_cyc_outc(const void* s, const void* m, const void* mx, int plus, const void* z, void* o, size_t n) {
    for (size_t i = 0; i < n; i++) { Fq r; cyc_outc(r, ld<Fq>(s, i), ld<Fq>(m, i), ld<Fq>(mx, i), plus != 0, ld<Fq>(z, i)); st(o, i, r); }
}

UN(g1_neg, G1, jac_neg) UN(g2_neg, G2, jac_neg)
extern "C" void hs_g1_eq(const void* a, const void* b, unsigned char* eq, size_t n) {
    for (size_t i = 0; i < n; i++) { G1 x = ld<G1>(a, i), y = ld<G1>(b, i); eq[i] = jac_eq(x, y); }
}
extern "C" void hs_g2_eq(const void* a, const void* b, unsigned char* eq, size_t n) {
    for (size_t i = 0; i < n; i++) { G2 x = ld<G2>(a, i), y = ld<G2>(b, i); eq[i] = jac_eq(x, y); }
}
extern "C" void hs_fq_inv(const void* a, void* o, unsigned char* some, size_t n) {
    for (size_t i = 0; i < n; i++) { Fq x = ld<Fq>(a, i), r; some[i] = fq_inv(r, x); st(o, i, r); }
}
extern "C" void hs_fq_inv_eea(const void* a, void* o, unsigned char* some, size_t n) {
    for (size_t i = 0; i < n; i++) { Fq x = ld<Fq>(a, i), r; some[i] = fq_inv_eea(r, x); st(o, i, r); }
}
extern "C" void hs_fq_inv_bgcd(const void* a, void* o, unsigned char* some, size_t n) {
    for (size_t i = 0; i < n; i++) { Fq x = ld<Fq>(a, i), r; some[i] = (unsigned char)fq_inv_bgcd_core(r, x); st(o, i, r); }   // 2 = fallback needed
}
extern "C" void hs_fq_inv_w4(const void* a, void* o, unsigned char* some, size_t n) {
    for (size_t i = 0; i < n; i++) { Fq x = ld<Fq>(a, i), r; some[i] = fq_inv_w4(r, x); st(o, i, r); }
}
extern "C" void hs_fq12_inv(const void* a, void* o, unsigned char* some, size_t n) {
    for (size_t i = 0; i < n; i++) { Fq12 x = ld<Fq12>(a, i), r; some[i] = fq12_inv(r, x); st(o, i, r); }
}
extern "C" void hs_final_exp(const void* a, void* o, unsigned char* some, size_t n) {
    for (size_t i = 0; i < n; i++) { Fq12 x = ld<Fq12>(a, i), r; some[i] = fq12_final_exponentiation(r, x); st(o, i, r); }
}
extern "C" void hs_fq12_frobenius(const void* a, int power, void* o, size_t n) {
    for (size_t i = 0; i < n; i++) { Fq12 x = ld<Fq12>(a, i), r; fq12_frobenius(r, x, power); st(o, i, r); }
}
extern "C" void hs_fq12_mul_by_024(const void* f, const void* ell, void* o, size_t n) {
    for (size_t i = 0; i < n; i++) { Fq12 x = ld<Fq12>(f, i); EllCoeffs c = ld<EllCoeffs>(ell, i); fq12_mul_by_024(x, c.ell_0, c.ell_vw, c.ell_vv); st(o, i, x); }
}
extern "C" void hs_gt_pow(const void* a, const void* fr, void* o, size_t n) {
    for (size_t i = 0; i < n; i++) {
        Fq12 x = ld<Fq12>(a, i), r; Fq m = ld<Fq>(fr, i); uint32_t k[8]; fr_to_canonical(k, m.v); fq12_pow(r, x, k); st(o, i, r);
    }
}
extern "C" void hs_g1_mul(const void* p, const void* fr, void* o, size_t n) {
    for (size_t i = 0; i < n; i++) { G1 x = ld<G1>(p, i), r; Fq m = ld<Fq>(fr, i); uint32_t k[8]; fr_to_canonical(k, m.v); jac_mul(r, x, k); st(o, i, r); }
}
extern "C" void hs_g2_mul(const void* p, const void* fr, void* o, size_t n) {
    for (size_t i = 0; i < n; i++) { G2 x = ld<G2>(p, i), r; Fq m = ld<Fq>(fr, i); uint32_t k[8]; fr_to_canonical(k, m.v); jac_mul(r, x, k); st(o, i, r); }
}
// (p * k).normalize() by the signed 4-bit window (k_jac_mul_window's body)
extern "C" void hs_g1_mul_normalized(const void* p, const void* fr, void* o, size_t n) {
    for (size_t i = 0; i < n; i++) { G1 x = ld<G1>(p, i), r; Fq m = ld<Fq>(fr, i); uint32_t k[8]; fr_to_canonical(k, m.v); jac_mul_normalized(r, x, k); st(o, i, r); }
}
extern "C" void hs_g2_mul_normalized(const void* p, const void* fr, void* o, size_t n) {
    for (size_t i = 0; i < n; i++) { G2 x = ld<G2>(p, i), r; Fq m = ld<Fq>(fr, i); uint32_t k[8]; fr_to_canonical(k, m.v); jac_mul_normalized(r, x, k); st(o, i, r); }
}
// digits of the recoding: out[i] = d_i in -7..8 for i < 64
extern "C" void hs_wnaf4_fixed_recode(const void* k_le, signed char* out) {
    uint32_t k[8], mag[8], sgn[2];
    memcpy(k, k_le, 32);
    wnaf4_fixed_recode(mag, sgn, k);
    for (int i = 0; i < 64; i++) {
        int a = (mag[i >> 3] >> (4 * (i & 7))) & 15;
        out[i] = (signed char)(((sgn[i >> 5] >> (i & 31)) & 1) ? -a : a);
    }
}
extern "C" void hs_fr_to_canonical(const void* fr, void* o, size_t n) {
    for (size_t i = 0; i < n; i++) { Fq m = ld<Fq>(fr, i), r; fr_to_canonical(r.v, m.v); st(o, i, r); }
}
// G2Precomp image: q (128 B) + 87 x EllCoeffs; returns 0 at infinity
extern "C" int hs_g2_precompute(const void* q, void* o) {
    G2 x = ld<G2>(q, 0); AffG2 qa;
    if (!jac_to_affine(qa, x)) return 0;
    EllCoeffs tab[BN_N_COEFFS];
    StoreSink s; s.out = tab; s.n = 0;
    g2_walk_lines(qa, s);
    if (s.n != BN_N_COEFFS) return -1;
    memcpy(o, &qa, sizeof qa);
    memcpy((char*)o + sizeof qa, tab, sizeof tab);
    return 1;
}
extern "C" void hs_miller_loop(const void* q, const void* p, void* o, unsigned char* ok, size_t n) {
    for (size_t i = 0; i < n; i++) {
        G1 pp = ld<G1>(p, i); G2 qq = ld<G2>(q, i); AffG1 pa; AffG2 qa; Fq12 f = fq12_one();
        bool good = jac_to_affine(pa, pp) & jac_to_affine(qa, qq);
        if (good) miller_loop_fused(f, pa, qa);
        ok[i] = good; st(o, i, f);
    }
}
extern "C" void hs_pairing(const void* p, const void* q, void* o, size_t n) {
    for (size_t i = 0; i < n; i++) { Fq12 r; pairing(r, ld<G1>(p, i), ld<G2>(q, i)); st(o, i, r); }
}
extern "C" void hs_constants(void* o) {
    char* p = (char*)o;
    Fq one = BN_C_ONE, r2 = BN_C_R2, r3 = BN_C_R3, two_inv = BN_C_TWO_INV, g1b = BN_C_G1_B, g1y = BN_C_G1_GEN_Y;
    Fq2 g2b = BN_C_G2_B, tx = BN_C_TWIST_MUL_BY_Q_X, ty = BN_C_TWIST_MUL_BY_Q_Y, gx = BN_C_G2_GEN_X, gy = BN_C_G2_GEN_Y;
    Fq2 f61[3] = {BN_C_FROB6_C1_1, BN_C_FROB6_C1_2, BN_C_FROB6_C1_3};
    Fq2 f62[3] = {BN_C_FROB6_C2_1, BN_C_FROB6_C2_2, BN_C_FROB6_C2_3};
    Fq2 f12[3] = {BN_C_FROB12_C1_1, BN_C_FROB12_C1_2, BN_C_FROB12_C1_3};
#define PUT(x) do { memcpy(p, &(x), sizeof(x)); p += sizeof(x); } while (0)
    PUT(one); PUT(r2); PUT(r3); PUT(two_inv); PUT(g1b); PUT(g1y); PUT(g2b); PUT(tx); PUT(ty);
    PUT(f61); PUT(f62); PUT(f12); PUT(gx); PUT(gy);
}

// validation, square roots, compressed wire formats (bn_wire.cuh)
extern "C" void hs_g1_affine_check(const void* xy, unsigned char* status, size_t n) {
    for (size_t i = 0; i < n; i++) status[i] = g1_affine_status(ld<AffG1>(xy, i));
}
extern "C" void hs_g2_affine_check(const void* xy, unsigned char* status, size_t n, int ref_ladder) {
    for (size_t i = 0; i < n; i++) status[i] = g2_affine_status(ld<AffG2>(xy, i), ref_ladder != 0);
}
extern "C" void hs_fq_sqrt(const void* a, void* o, unsigned char* some, size_t n) {
    for (size_t i = 0; i < n; i++) { Fq x = ld<Fq>(a, i), r = fq_zero(); some[i] = fq_sqrt(r, x); st(o, i, r); }
}
extern "C" void hs_fq2_sqrt(const void* a, void* o, unsigned char* some, size_t n) {
    for (size_t i = 0; i < n; i++) { Fq2 x = ld<Fq2>(a, i), r = fq2_zero(); some[i] = fq2_sqrt(r, x); st(o, i, r); }
}
extern "C" void hs_fq2_sqrt_any(const void* a, void* o, unsigned char* some, size_t n) {
    for (size_t i = 0; i < n; i++) { Fq2 x = ld<Fq2>(a, i), r = fq2_zero(); some[i] = fq2_sqrt_any(r, x); st(o, i, r); }
}
extern "C" void hs_g1_from_compressed(const unsigned char* enc, void* o, unsigned char* status, size_t n) {
    for (size_t i = 0; i < n; i++) {
        G1 g; memset(&g, 0, sizeof g);
        G1 t;
        status[i] = g1_decode(t, enc + 33 * i);
        if (status[i] == 0) g = t;
        st(o, i, g);
    }
}
extern "C" void hs_g2_from_compressed(const unsigned char* enc, void* o, unsigned char* status, size_t n, int ref_ladder) {
    for (size_t i = 0; i < n; i++) {
        G2 g; memset(&g, 0, sizeof g);
        G2 t;
        status[i] = g2_decode(t, enc + 65 * i, ref_ladder != 0);
        if (status[i] == 0) g = t;
        st(o, i, g);
    }
}
extern "C" int hs_u512_divrem_q(const unsigned char* be, void* quot, void* rem) {
    uint32_t q[8], r[8];
    bool ok = u512_divrem_q(q, r, be);
    memcpy(quot, q, 32);
    memcpy(rem, r, 32);
    return ok ? 1 : 0;
}
