// G2 line steps on the shared-memory slot file (device only): AffineG2::precompute, src/groups/mod.rs:701-776.
//
// The running point R = (X, Y, Z) stays in slots 0..2 for the whole walk; slots 3..6 are scratch.  Products run through
// s_mul / s_sqr, the additive work between them in a few register-resident stages that also store the line coefficients
// straight into the word-major table.  Same field elements as g2_doubling_step / g2_mixed_addition_step (bn_curve.cuh).
#pragma once
#include "bn_curve.cuh"
#include "bn_slots.cuh"

#if defined(__CUDACC__)
namespace bn {

__device__ __noinline__ void s_sqr(sp r, sp a) { Fq2 x = s_ld(a), z; fq2_sqr_inl(z, x); s_st(r, z); }

// a/2 on registers (fq_half of bn_curve.cuh goes through memory)
__device__ __forceinline__ void fq_half_r(Fq& r, const Fq& a) {
    uint32_t mask = 0u - (a.v[0] & 1u);
    uint32_t t[9];
    uint64_t c = 0;
#pragma unroll
    for (int i = 0; i < 8; i++) {
        uint64_t s = (uint64_t)a.v[i] + (fq_modulus_limb(i) & mask) + c;
        t[i] = (uint32_t)s;
        c = s >> 32;
    }
    t[8] = (uint32_t)c;
#pragma unroll
    for (int i = 0; i < 8; i++) r.v[i] = (t[i] >> 1) | (t[i + 1] << 31);
}
__device__ __forceinline__ void fq2_add_r(Fq2& r, const Fq2& a, const Fq2& b) { fq_add(r.c0, a.c0, b.c0); fq_add(r.c1, a.c1, b.c1); }
__device__ __forceinline__ void fq2_sub_r(Fq2& r, const Fq2& a, const Fq2& b) { fq_sub(r.c0, a.c0, b.c0); fq_sub(r.c1, a.c1, b.c1); }
__device__ __forceinline__ void fq2_neg_r(Fq2& r, const Fq2& a) { fq_neg(r.c0, a.c0); fq_neg(r.c1, a.c1); }
__device__ __forceinline__ void fq2_xi_r(Fq2& r, const Fq2& a) {
    Fq2 n;
    fq_lin9<true>(n.c0, a.c0, a.c1);
    fq_lin9<false>(n.c1, a.c1, a.c0);
    r = n;
}
// one Fq2 of a coefficient triple into the word-major table: words w0..w0+3 of triple idx (layout of load_coeff)
struct LineOut {
    uint4* tab;        // table base + column of this thread's table
    size_t stride;     // number of tables (words of one position are `stride` apart)
    bool live;         // false: dummy thread, nothing is stored
};
__device__ __forceinline__ void line_store(const LineOut& o, int idx, int w0, const Fq2& v) {
    if (!o.live) return;
    uint4* p = o.tab + ((size_t)idx * 12 + w0) * o.stride;
    p[0] = make_uint4(v.c0.v[0], v.c0.v[1], v.c0.v[2], v.c0.v[3]);
    p[o.stride] = make_uint4(v.c0.v[4], v.c0.v[5], v.c0.v[6], v.c0.v[7]);
    p[2 * o.stride] = make_uint4(v.c1.v[0], v.c1.v[1], v.c1.v[2], v.c1.v[3]);
    p[3 * o.stride] = make_uint4(v.c1.v[4], v.c1.v[5], v.c1.v[6], v.c1.v[7]);
}

// ---- doubling step (src/groups/mod.rs:754-776) ---------------------------------------------------------------------
// slots on entry/exit: 0 = X, 1 = Y, 2 = Z.  10 products (4 s_mul + 6 s_sqr), 4 additive stages.
// stage 2: with j = X^2 (0), X Y (3), (Y+Z)^2 (1), b = Y^2 (4), c = Z^2 (2):
//   ell_vv = 3 j, a = X Y / 2 -> 3, h = (Y+Z)^2 - b - c -> 1, ell_vw = -h, 5 <- 3 b'
__device__ __noinline__ void s_dbl_g2(sp base, LineOut o, int idx) {
    const Fq2 b3 = BN_C_G2_B3;
    Fq2 x = s_ld(base), t;
    fq2_add_r(t, x, x); fq2_add_r(t, t, x);
    line_store(o, idx, 8, t);                                  // ell_vv = 3 X^2
    x = s_ld(base + 192);
    fq_half_r(t.c0, x.c0); fq_half_r(t.c1, x.c1);
    s_st(base + 192, t);                                       // a
    x = s_ld(base + 64);
    t = s_ld(base + 256); fq2_sub_r(x, x, t);
    t = s_ld(base + 128); fq2_sub_r(x, x, t);
    s_st(base + 64, x);                                        // h
    fq2_neg_r(t, x);
    line_store(o, idx, 4, t);                                  // ell_vw = -h
    s_st(base + 320, b3);
}
// stage 3: with e (2), b (4): ell_0 = xi (e - b), f = 3 e, g = (b + f) / 2 -> 6, b - f -> 0
__device__ __noinline__ void s_dbl_g3(sp base, LineOut o, int idx) {
    Fq2 e = s_ld(base + 128), b = s_ld(base + 256), t, f;
    fq2_sub_r(t, e, b);
    fq2_xi_r(t, t);
    line_store(o, idx, 0, t);                                  // ell_0
    fq2_add_r(f, e, e); fq2_add_r(f, f, e);
    fq2_add_r(t, b, f);
    fq_half_r(e.c0, t.c0); fq_half_r(e.c1, t.c1);
    s_st(base + 384, e);                                       // g
    fq2_sub_r(t, b, f);
    s_st(base, t);
}
// stage 4: Y' = g^2 (6) - 3 e^2 (5) -> 1
__device__ __noinline__ void s_dbl_g4(sp base) {
    Fq2 g = s_ld(base + 384), q = s_ld(base + 320), t;
    fq2_add_r(t, q, q); fq2_add_r(t, t, q);
    fq2_sub_r(g, g, t);
    s_st(base + 64, g);
}
__device__ __noinline__ void s_g2_doubling_step(sp base, LineOut o, int idx) {
    const sp S0 = base, S1 = base + 64, S2 = base + 128, S3 = base + 192, S4 = base + 256, S5 = base + 320, S6 = base + 384;
    s_mul(S3, S0, S1);            // X Y
    s_sqr(S0, S0);                // j = X^2
    s_sqr(S4, S1);                // b = Y^2
    s_add(S1, S1, S2);            // Y + Z
    s_sqr(S2, S2);                // c = Z^2
    s_sqr(S1, S1);                // (Y + Z)^2
    s_dbl_g2(base, o, idx);
    s_mul(S2, S5, S2);            // e = 3 b' c
    s_dbl_g3(base, o, idx);
    s_sqr(S5, S2);                // e^2
    s_mul(S0, S3, S0);            // X' = a (b - f)
    s_sqr(S6, S6);                // g^2
    s_mul(S2, S4, S1);            // Z' = b h
    s_dbl_g4(base);
}

// ---- mixed addition step (src/groups/mod.rs:731-752) ----------------------------------------------------------------
// 11 s_mul + 2 s_sqr; the affine point (x, y) is read from local memory where a product needs it.
// stage 1: ell_vw = d (3), ell_vv = -e (4)
__device__ __noinline__ void s_add_g1(sp base, LineOut o, int idx) {
    Fq2 d = s_ld(base + 192), e = s_ld(base + 256), t;
    line_store(o, idx, 4, d);
    fq2_neg_r(t, e);
    line_store(o, idx, 8, t);
}
// stage 2: j = Z g (6) + h (5) - 2 i (0) -> 6, i - j -> 0
__device__ __noinline__ void s_add_g2(sp base) {
    Fq2 j = s_ld(base + 384), h = s_ld(base + 320), i = s_ld(base), t;
    fq2_add_r(j, j, h);
    fq2_add_r(t, i, i);
    fq2_sub_r(j, j, t);
    s_st(base + 384, j);
    fq2_sub_r(t, i, j);
    s_st(base, t);
}
// stage 3: ell_0 = xi (e x (5) - d y (6))
__device__ __noinline__ void s_add_g3(sp base, LineOut o, int idx) {
    Fq2 a = s_ld(base + 320), b = s_ld(base + 384);
    fq2_sub_r(a, a, b);
    fq2_xi_r(a, a);
    line_store(o, idx, 0, a);
}
__device__ __noinline__ void s_g2_mixed_addition_step(sp base, const Fq2& qx, const Fq2& qy, LineOut o, int idx) {
    const sp S0 = base, S1 = base + 64, S2 = base + 128, S3 = base + 192, S4 = base + 256, S5 = base + 320, S6 = base + 384;
    s_in(S3, qx); s_mul(S3, S2, S3); s_sub(S3, S0, S3);     // d = X - Z x
    s_in(S4, qy); s_mul(S4, S2, S4); s_sub(S4, S1, S4);     // e = Y - Z y
    s_add_g1(base, o, idx);
    s_sqr(S5, S3);                // f = d^2
    s_sqr(S6, S4);                // g = e^2
    s_mul(S0, S0, S5);            // i = X f
    s_mul(S5, S3, S5);            // h = d f
    s_mul(S6, S2, S6);            // Z g
    s_add_g2(base);
    s_mul(S0, S4, S0);            // e (i - j)
    s_mul(S1, S5, S1);            // h Y
    s_sub(S1, S0, S1);            // Y'
    s_mul(S2, S2, S5);            // Z' = Z h
    s_mul(S0, S3, S6);            // X' = d j
    s_in(S5, qx); s_mul(S5, S4, S5);                        // e x
    s_in(S6, qy); s_mul(S6, S3, S6);                        // d y
    s_add_g3(base, o, idx);
}

}  // namespace bn
#endif
