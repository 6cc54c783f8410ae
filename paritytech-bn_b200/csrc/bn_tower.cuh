// Field tower Fq2 -> Fq6 -> Fq12 for BN254, "lean" evaluation of the reference's formulas:
// same field elements (hence the same fully-reduced Montgomery bytes) as src/fields/{fq2,fq6,fq12}.rs,
// but multiplication by -1, by xi = 9+u and conjugation are done with additions, Fq2 products use
// lazy reduction (3 wide products, 2 Montgomery reductions), and inversions use Fermat (a^(q-2)) --
// SURVEY.md 8a rows A3, A5-A8, A14-A16 explain why any exact evaluation is bit-identical.
//
// All functions are alias-safe (outputs may alias inputs).  Big ones are __noinline__ on the device
// so the kernels stay a few hundred KB of SASS and Fq12 state lives in per-thread local memory
// while each Fq2-level body is register-blocked.
#pragma once
#include "bn_consts.cuh"
#include "bn_fq.cuh"

namespace bn {

// Fq2-level additive helpers are CALLS on the device by default: measured on B200 the Miller and
// final-exponentiation kernels are instruction-fetch bound (ncu: stall "no_instruction" on top), and
// out-of-line adds shrink k_miller_many from 180 KB to 76 KB of SASS (+15% pairings/s at 128
// registers).  -DBN_INLINE_ADDS restores inlining (kept for A/B runs).
#if defined(__CUDACC__) && !defined(BN_INLINE_ADDS)
#define BN_ADD_FN BN_HD_NOINLINE
#else
#define BN_ADD_FN BN_HD
#endif

struct Fq2 { Fq c0, c1; };            // src/fields/fq2.rs:36-41 (repr(C))
struct Fq6 { Fq2 c0, c1, c2; };       // src/fields/fq6.rs:92-98
struct Fq12 { Fq6 c0, c1; };          // src/fields/fq12.rs:50-55

BN_HD Fq fq_one() { Fq r = BN_C_ONE; return r; }
BN_HD Fq fq_zero() { Fq r; fq_set_zero(r); return r; }

// ---------------------------------------------------------------------------------------------
// Fq helpers: Fermat inversion (replaces the divergent binary EEA of src/arith.rs:324-370)
// ---------------------------------------------------------------------------------------------
BN_HD_NOINLINE void fq_mul_ni(Fq& r, const Fq& a, const Fq& b) { Fq x = fq_ld(a), y = fq_ld(b), z; fq_mul(z, x, y); fq_st(r, z); }

// r = a^(q-2); returns false (and r = 0) for a == 0, like Fq::inverse -> None (src/fields/fp.rs:108-117)
BN_HD_NOINLINE bool fq_inv(Fq& r, const Fq& a) {
    if (fq_is_zero(a)) { fq_set_zero(r); return false; }
    const uint32_t e[8] = BN_Q_MINUS_2_LIMBS;
    Fq acc = a;  // top bit (bit 253) is set
    for (int i = 252; i >= 0; i--) {
        fq_mul_ni(acc, acc, acc);
        if ((e[i >> 5] >> (i & 31)) & 1) fq_mul_ni(acc, acc, a);
    }
    r = acc;
    return true;
}
// Same value with a fixed 4-bit window (14 + 252 + 59 multiplications instead of 253 + 109): used where the
// inversion is a visible share of the kernel (k_pairs_to_affine: -10 %).  Kept out of the final-exponentiation
// kernel, whose single inversion is 1 % of its work and whose code layout the larger body disturbed (+1 %).
// q - 2 has 254 bits = a 2-bit top window (0b11) followed by 63 nibbles.
BN_HD_NOINLINE bool fq_inv_w4(Fq& r, const Fq& a) {
    if (fq_is_zero(a)) { fq_set_zero(r); return false; }
    const uint32_t e[8] = BN_Q_MINUS_2_LIMBS;
    Fq tab[16];
    tab[1] = a;
    for (int i = 2; i < 16; i++) fq_mul_ni(tab[i], tab[i - 1], a);
    Fq acc = tab[3];
    for (int w = 62; w >= 0; w--) {
        fq_mul_ni(acc, acc, acc); fq_mul_ni(acc, acc, acc); fq_mul_ni(acc, acc, acc); fq_mul_ni(acc, acc, acc);
        uint32_t nib = (e[w >> 3] >> ((w & 7) * 4)) & 15u;
        if (nib) fq_mul_ni(acc, acc, tab[nib]);
    }
    r = acc;
    return true;
}

// Binary extended Euclid, the reference's own method (U256::invert, src/arith.rs:324-370; Fq::inverse src/fields/fp.rs:108-117)
// rearranged for SIMT: ONE uniform step per iteration, no data-dependent branches inside the loop --
//     if u is odd: make u the larger of the two (conditional swap of (u, b) with (v, c)), then u -= v, b -= c (mod q);
//     u /= 2, b /= 2 (mod q)
// with the invariants b x = u, c x = v (mod q), v odd.  bitlen(u) + bitlen(v) drops by at least one per step, so the loop ends
// within 508 steps with u = 0, v = 1, c = x^-1.  x is the Montgomery residue a R, so c = a^-1 R^-1 and one Montgomery
// multiplication by R^3 gives a^-1 R -- exactly the reference's sequence.  All ALU work (~150 instructions per step): a third of
// the multiplier-pipe time of the Fermat power (325-362 Montgomery multiplications); the same field element, hence the same bytes.
BN_HD_NOINLINE bool fq_inv_eea(Fq& r, const Fq& a) {
    if (fq_is_zero(a)) { fq_set_zero(r); return false; }
    uint32_t u[8], v[8], b[8], c[8];
#pragma unroll
    for (int i = 0; i < 8; i++) { u[i] = a.v[i]; v[i] = fq_modulus_limb(i); b[i] = 0; c[i] = 0; }
    b[0] = 1;
    uint32_t nz = 1;
#pragma unroll 1
    while (nz) {
        const uint32_t odd = 0u - (u[0] & 1u);
        // lt = (u < v): borrow out of u - v
        uint32_t bw = 0;
#pragma unroll
        for (int i = 0; i < 8; i++) {
            uint64_t d = (uint64_t)u[i] - v[i] - bw;
            bw = (uint32_t)(d >> 63);
        }
        const uint32_t sw = odd & (0u - bw);          // swap when u is odd and smaller than v
#pragma unroll
        for (int i = 0; i < 8; i++) {
            uint32_t t = (u[i] ^ v[i]) & sw; u[i] ^= t; v[i] ^= t;
            t = (b[i] ^ c[i]) & sw; b[i] ^= t; c[i] ^= t;
        }
        // u -= v, b -= c (mod q) when u was odd
        bw = 0;
#pragma unroll
        for (int i = 0; i < 8; i++) {
            uint64_t d = (uint64_t)u[i] - (v[i] & odd) - bw;
            u[i] = (uint32_t)d; bw = (uint32_t)(d >> 63);
        }
        bw = 0;
#pragma unroll
        for (int i = 0; i < 8; i++) {
            uint64_t d = (uint64_t)b[i] - (c[i] & odd) - bw;
            b[i] = (uint32_t)d; bw = (uint32_t)(d >> 63);
        }
        uint32_t fix = 0u - bw, cy = 0;               // b < c: add q back
#pragma unroll
        for (int i = 0; i < 8; i++) {
            uint64_t t = (uint64_t)b[i] + (fq_modulus_limb(i) & fix) + cy;
            b[i] = (uint32_t)t; cy = (uint32_t)(t >> 32);
        }
        // u /= 2 (u is even now); b /= 2 mod q: (b + (b odd ? q : 0)) >> 1, the sum below 2^255
        const uint32_t bodd = 0u - (b[0] & 1u);
        cy = 0;
#pragma unroll
        for (int i = 0; i < 8; i++) {
            uint64_t t = (uint64_t)b[i] + (fq_modulus_limb(i) & bodd) + cy;
            b[i] = (uint32_t)t; cy = (uint32_t)(t >> 32);
        }
        nz = 0;
#pragma unroll
        for (int i = 0; i < 7; i++) {
            u[i] = (u[i] >> 1) | (u[i + 1] << 31);
            b[i] = (b[i] >> 1) | (b[i + 1] << 31);
            nz |= u[i];
        }
        u[7] >>= 1; b[7] >>= 1;
        nz |= u[7];
    }
    Fq inv, r3 = BN_C_R3;
#pragma unroll
    for (int i = 0; i < 8; i++) inv.v[i] = c[i];
    fq_mul_ni(r, inv, r3);
    return true;
}

// The same binary extended Euclid in rounds of 30 steps on 64-bit APPROXIMATIONS of (u, v) -- the low 30 bits (which decide the
// parities of 30 steps exactly) under the top 34 bits of the common bit window (which decide the comparisons) -- collecting the
// steps in a 2 x 2 integer matrix (entries |f| + |g| <= 2^30) that is then applied once to the full-size (u, v) and, modulo q
// with an exact division by 2^30 (Montgomery style, q' = -q^-1 mod 2^30), to (b, c).  A comparison decided on the approximation
// can be wrong only when the two values agree in their top 34 bits; the update then comes out negative and is negated together
// with its matrix row.  17 rounds = 510 steps >= 2 * 254 - 1.  ~1,250 instructions per round against ~6,000 for 30 plain steps;
// the result is CHECKED (u = 0, v = 1) and the plain loop above is the fallback, so the value never depends on the
// approximation argument.  Same field element, hence the same bytes.  Python model: tools/bgcd_model.py.
#define BN_BGCD_K 30
#define BN_BGCD_ROUNDS 17
#define BN_BGCD_NEGQINV 0x24866389u                  /* -q^-1 mod 2^30 */
// out[9] = f x + g y as a 288-bit two's-complement integer (|f| + |g| <= 2^30, x, y < 2^256)
BN_HD void bgcd_lin2(uint32_t (&out)[9], int32_t f, const uint32_t (&x)[8], int32_t g, const uint32_t (&y)[8]) {
    long long acc = 0;
#pragma unroll
    for (int i = 0; i < 8; i++) {
        acc += (long long)f * (long long)(unsigned long long)x[i] + (long long)g * (long long)(unsigned long long)y[i];
        out[i] = (uint32_t)acc;
        acc >>= 32;
    }
    out[8] = (uint32_t)acc;
}
BN_HD uint32_t bgcd_shl_hi(uint32_t hi, uint32_t lo, uint32_t s) { return s ? (hi << s) | (lo >> (32u - s)) : hi; }
// returns 0: x is zero (r = 0), 1: r = the inverse, 2: the rounds did not end in (0, 1) (r untouched)
BN_HD_NOINLINE int fq_inv_bgcd_core(Fq& r, const Fq& x) {
    if (fq_is_zero(x)) { fq_set_zero(r); return 0; }
    uint32_t a[8], b[8], u[8], v[8];
#pragma unroll
    for (int i = 0; i < 8; i++) { a[i] = x.v[i]; b[i] = fq_modulus_limb(i); u[i] = 0; v[i] = 0; }
    u[0] = 1;
    const uint32_t lowmask = (1u << BN_BGCD_K) - 1u;
#pragma unroll 1
    for (int round = 0; round < BN_BGCD_ROUNDS; round++) {
        // the three limbs under the top non-zero limb of a | b
        uint32_t ah = a[1], am = a[0], al = 0, bh = b[1], bm = b[0], bl = 0, ch = 1, big = 0;
#pragma unroll
        for (int i = 2; i < 8; i++) {
            const uint32_t c = a[i] | b[i];
            const bool nz = c != 0;
            ah = nz ? a[i] : ah; am = nz ? a[i - 1] : am; al = nz ? a[i - 2] : al;
            bh = nz ? b[i] : bh; bm = nz ? b[i - 1] : bm; bl = nz ? b[i - 2] : bl;
            ch = nz ? c : ch; big = nz ? 1u : big;
        }
#if defined(__CUDA_ARCH__)
        const uint32_t sh = (uint32_t)__clz((int)ch);
#else
        uint32_t sh = 0;
        for (uint32_t t = ch; !(t & 0x80000000u); t <<= 1) sh++;                 // count of leading zeros of the top limb
#endif
        unsigned long long a_ = ((unsigned long long)bgcd_shl_hi(ah, am, sh) << 32) | bgcd_shl_hi(am, al, sh);
        unsigned long long b_ = ((unsigned long long)bgcd_shl_hi(bh, bm, sh) << 32) | bgcd_shl_hi(bm, bl, sh);
        a_ = ((a_ >> BN_BGCD_K) << BN_BGCD_K) | (a[0] & lowmask);
        b_ = ((b_ >> BN_BGCD_K) << BN_BGCD_K) | (b[0] & lowmask);
        if (!big) {                                                             // both fit 64 bits: exact
            a_ = ((unsigned long long)a[1] << 32) | a[0];
            b_ = ((unsigned long long)b[1] << 32) | b[0];
        }
        uint32_t f0 = 1, g0 = 0, f1 = 0, g1 = 1;                                // two's complement
#pragma unroll 2
        for (int it = 0; it < BN_BGCD_K; it++) {
            const unsigned long long odd = 0ull - (a_ & 1ull);
            const unsigned long long sw = (a_ < b_ ? ~0ull : 0ull) & odd;
            const unsigned long long t = (a_ ^ b_) & sw;
            a_ ^= t; b_ ^= t;
            const uint32_t sw32 = (uint32_t)sw, odd32 = (uint32_t)odd;
            uint32_t tf = (f0 ^ f1) & sw32; f0 ^= tf; f1 ^= tf;
            tf = (g0 ^ g1) & sw32; g0 ^= tf; g1 ^= tf;
            a_ -= b_ & odd; f0 -= f1 & odd32; g0 -= g1 & odd32;
            a_ >>= 1; f1 <<= 1; g1 <<= 1;
        }
        // (a, b) <- (f0 a + g0 b, f1 a + g1 b) / 2^30, negated together with their rows when negative
        uint32_t na[9], nb[9];
        bgcd_lin2(na, (int32_t)f0, a, (int32_t)g0, b);
        bgcd_lin2(nb, (int32_t)f1, a, (int32_t)g1, b);
        const uint32_t sa = 0u - (na[8] >> 31), sb = 0u - (nb[8] >> 31);
        uint32_t ca = sa & 1u, cb = sb & 1u;
#pragma unroll
        for (int i = 0; i < 8; i++) {
            uint32_t wa = ((na[i] >> BN_BGCD_K) | (na[i + 1] << (32 - BN_BGCD_K))) ^ sa;
            uint32_t wb = ((nb[i] >> BN_BGCD_K) | (nb[i + 1] << (32 - BN_BGCD_K))) ^ sb;
            unsigned long long ta = (unsigned long long)wa + ca, tb = (unsigned long long)wb + cb;
            a[i] = (uint32_t)ta; ca = (uint32_t)(ta >> 32);
            b[i] = (uint32_t)tb; cb = (uint32_t)(tb >> 32);
        }
        f0 = (f0 ^ sa) - sa; g0 = (g0 ^ sa) - sa; f1 = (f1 ^ sb) - sb; g1 = (g1 ^ sb) - sb;
        // (u, v) <- the same rows applied modulo q, divided exactly by 2^30
        uint32_t nu[9], nv[9];
        bgcd_lin2(nu, (int32_t)f0, u, (int32_t)g0, v);
        bgcd_lin2(nv, (int32_t)f1, u, (int32_t)g1, v);
#pragma unroll
        for (int half = 0; half < 2; half++) {
            uint32_t (&t)[9] = half ? nv : nu;
            uint32_t (&o)[8] = half ? v : u;
            const uint32_t m = ((t[0] & lowmask) * BN_BGCD_NEGQINV) & lowmask;
            unsigned long long cy = 0;
#pragma unroll
            for (int i = 0; i < 8; i++) {
                cy += (unsigned long long)m * fq_modulus_limb(i) + t[i];
                t[i] = (uint32_t)cy; cy >>= 32;
            }
            t[8] += (uint32_t)cy;
            const uint32_t neg = 0u - (t[8] >> 31);                              // the quotient lies in (-q, 2q)
            uint32_t w[8], p[8], mns[8];
            unsigned long long c1 = 0;
            long long br = 0;
#pragma unroll
            for (int i = 0; i < 8; i++) {
                w[i] = (t[i] >> BN_BGCD_K) | (t[i + 1] << (32 - BN_BGCD_K));
                c1 += (unsigned long long)w[i] + fq_modulus_limb(i);
                p[i] = (uint32_t)c1; c1 >>= 32;
                br += (long long)w[i] - (long long)fq_modulus_limb(i);
                mns[i] = (uint32_t)br; br >>= 32;
            }
            const uint32_t lt = (uint32_t)br;                                    // all ones when w < q
#pragma unroll
            for (int i = 0; i < 8; i++) o[i] = neg ? p[i] : (lt ? w[i] : mns[i]);
        }
    }
    uint32_t rest = b[0] ^ 1u;
#pragma unroll
    for (int i = 0; i < 8; i++) rest |= a[i] | (i ? b[i] : 0u);
    if (rest) return 2;
    Fq inv, r3 = BN_C_R3;
#pragma unroll
    for (int i = 0; i < 8; i++) inv.v[i] = v[i];
    fq_mul_ni(r, inv, r3);
    return 1;
}
BN_HD bool fq_inv_bgcd(Fq& r, const Fq& x) {
    const int st = fq_inv_bgcd_core(r, x);
    if (st == 2) return fq_inv_eea(r, x);                                        // never seen; keeps the value independent of the bound
    return st == 1;
}

// ---------------------------------------------------------------------------------------------
// Fq2 = Fq[u]/(u^2+1)
// ---------------------------------------------------------------------------------------------
BN_HD Fq2 fq2_zero() { Fq2 r; fq_set_zero(r.c0); fq_set_zero(r.c1); return r; }
BN_HD Fq2 fq2_one() { Fq2 r; r.c0 = fq_one(); fq_set_zero(r.c1); return r; }
BN_HD bool fq2_is_zero(const Fq2& a) { return fq_is_zero(a.c0) && fq_is_zero(a.c1); }
BN_HD bool fq2_eq(const Fq2& a, const Fq2& b) { return fq_eq(a.c0, b.c0) && fq_eq(a.c1, b.c1); }
BN_HD Fq2 fq2_ld(const Fq2& p) { Fq2 r; r.c0 = fq_ld(p.c0); r.c1 = fq_ld(p.c1); return r; }
BN_HD void fq2_st(Fq2& p, const Fq2& v) { fq_st(p.c0, v.c0); fq_st(p.c1, v.c1); }
// memory-to-memory copies of tower elements with 128-bit accesses (alias-safe for d == s)
template <class T> BN_HD void cpv(T& d, const T& s) {
#if defined(__CUDA_ARCH__)
    const uint4* sp = reinterpret_cast<const uint4*>(&s);
    uint4* dp = reinterpret_cast<uint4*>(&d);
    uint4 t[sizeof(T) / 16];
#pragma unroll
    for (int i = 0; i < (int)(sizeof(T) / 16); i++) t[i] = sp[i];
#pragma unroll
    for (int i = 0; i < (int)(sizeof(T) / 16); i++) dp[i] = t[i];
#else
    d = s;
#endif
}
BN_ADD_FN void fq2_add(Fq2& r, const Fq2& a, const Fq2& b) {
    Fq2 x = fq2_ld(a), y = fq2_ld(b), z;
    fq_add(z.c0, x.c0, y.c0); fq_add(z.c1, x.c1, y.c1);
    fq2_st(r, z);
}
BN_ADD_FN void fq2_sub(Fq2& r, const Fq2& a, const Fq2& b) {
    Fq2 x = fq2_ld(a), y = fq2_ld(b), z;
    fq_sub(z.c0, x.c0, y.c0); fq_sub(z.c1, x.c1, y.c1);
    fq2_st(r, z);
}
BN_ADD_FN void fq2_neg(Fq2& r, const Fq2& a) { Fq2 x = fq2_ld(a), z; fq_neg(z.c0, x.c0); fq_neg(z.c1, x.c1); fq2_st(r, z); }
BN_ADD_FN void fq2_dbl(Fq2& r, const Fq2& a) { Fq2 x = fq2_ld(a), z; fq_dbl(z.c0, x.c0); fq_dbl(z.c1, x.c1); fq2_st(r, z); }
BN_HD void fq2_conj(Fq2& r, const Fq2& a) { Fq2 x = fq2_ld(a); fq_neg(x.c1, x.c1); fq2_st(r, x); }   // frobenius_map(odd), fq2.rs:59-68

// Karatsuba with lazy reduction; value of src/fields/fq2.rs:133-149
BN_HD void fq2_mul_inl(Fq2& r, const Fq2& a, const Fq2& b) {
    FqWide t0, t1, s;
    Fq sa, sb;
    fq_mulwide(t0, a.c0, b.c0);
    fq_mulwide(t1, a.c1, b.c1);
    fq_add_noreduce(sa, a.c0, a.c1);
    fq_add_noreduce(sb, b.c0, b.c1);
    fq_mulwide(s, sa, sb);          // < 4q^2 < q*2^256
    wide_sub_nb(s, t0);                // >= 0: a0*b1 + a1*b0 + t1
    wide_sub_nb(s, t1);                // a0*b1 + a1*b0 < 2q^2
    wide_sub(t0, t1);               // a0*b0 - a1*b1 (+ q*2^256 if negative)
    fq_redc(r.c0, t0);
    fq_redc(r.c1, s);
}
BN_HD_NOINLINE void fq2_mul(Fq2& r, const Fq2& a, const Fq2& b) { Fq2 x = fq2_ld(a), y = fq2_ld(b), z; fq2_mul_inl(z, x, y); fq2_st(r, z); }

// complex squaring; value of src/fields/fq2.rs:105-117
BN_HD void fq2_sqr_inl(Fq2& r, const Fq2& a) {
    Fq s, d, m;
    fq_add_noreduce(s, a.c0, a.c1);   // < 2q, fine as a Montgomery multiplicand
    fq_sub(d, a.c0, a.c1);
    fq_mul(m, a.c0, a.c1);
    fq_mul(r.c0, s, d);
    fq_dbl(r.c1, m);
}
BN_HD_NOINLINE void fq2_sqr(Fq2& r, const Fq2& a) { Fq2 x = fq2_ld(a), z; fq2_sqr_inl(z, x); fq2_st(r, z); }

BN_HD_NOINLINE void fq2_scale(Fq2& r, const Fq2& a, const Fq& k) {   // src/fields/fq2.rs:48-53
    Fq2 x = fq2_ld(a), z;
    Fq kk = fq_ld(k);
    fq_mul(z.c0, x.c0, kk);
    fq_mul(z.c1, x.c1, kk);
    fq2_st(r, z);
}
// a * (9 + u) = (9 a0 - a1) + (9 a1 + a0) u; value of src/fields/fq2.rs:55-57
BN_ADD_FN void fq2_mul_by_xi(Fq2& r, const Fq2& aa) {
    Fq2 a = fq2_ld(aa), n;
    fq_lin9<true>(n.c0, a.c0, a.c1);      // 9 a0 - a1
    fq_lin9<false>(n.c1, a.c1, a.c0);     // 9 a1 + a0
    fq2_st(r, n);
}
BN_HD_NOINLINE bool fq2_inv(Fq2& r, const Fq2& a) {   // src/fields/fq2.rs:119-130
    Fq n, t, t2;
    fq_sqr(n, a.c0);
    fq_sqr(t, a.c1);
    fq_add(n, n, t);
    if (!fq_inv_bgcd(t, n)) { r = fq2_zero(); return false; }
    fq_mul(r.c0, a.c0, t);
    fq_mul(t2, a.c1, t);
    fq_neg(r.c1, t2);
    return true;
}

// ---------------------------------------------------------------------------------------------
// Fq6 = Fq2[v]/(v^3 - xi)
// ---------------------------------------------------------------------------------------------
BN_HD Fq6 fq6_zero() { Fq6 r; r.c0 = fq2_zero(); r.c1 = fq2_zero(); r.c2 = fq2_zero(); return r; }
BN_HD Fq6 fq6_one() { Fq6 r = fq6_zero(); r.c0.c0 = fq_one(); return r; }
BN_HD void fq6_add(Fq6& r, const Fq6& a, const Fq6& b) { fq2_add(r.c0, a.c0, b.c0); fq2_add(r.c1, a.c1, b.c1); fq2_add(r.c2, a.c2, b.c2); }
BN_HD void fq6_sub(Fq6& r, const Fq6& a, const Fq6& b) { fq2_sub(r.c0, a.c0, b.c0); fq2_sub(r.c1, a.c1, b.c1); fq2_sub(r.c2, a.c2, b.c2); }
BN_HD void fq6_neg(Fq6& r, const Fq6& a) { fq2_neg(r.c0, a.c0); fq2_neg(r.c1, a.c1); fq2_neg(r.c2, a.c2); }
BN_HD void fq6_dbl(Fq6& r, const Fq6& a) { fq2_dbl(r.c0, a.c0); fq2_dbl(r.c1, a.c1); fq2_dbl(r.c2, a.c2); }
BN_HD bool fq6_is_zero(const Fq6& a) { return fq2_is_zero(a.c0) && fq2_is_zero(a.c1) && fq2_is_zero(a.c2); }
BN_HD void fq6_mul_by_v(Fq6& r, const Fq6& a) {   // mul_by_nonresidue, src/fields/fq6.rs:109-115
    Fq2 t;
    fq2_mul_by_xi(t, a.c2);
    cpv(r.c2, a.c1);
    cpv(r.c1, a.c0);
    cpv(r.c0, t);
}
BN_HD_NOINLINE void fq6_mul(Fq6& r, const Fq6& a, const Fq6& b) {   // value of src/fields/fq6.rs:194-208
    Fq2 aa, bb, cc, t, u, s;
    fq2_mul(aa, a.c0, b.c0);
    fq2_mul(bb, a.c1, b.c1);
    fq2_mul(cc, a.c2, b.c2);
    Fq2 n0, n1, n2;
    fq2_add(t, a.c1, a.c2); fq2_add(u, b.c1, b.c2); fq2_mul(s, t, u);
    fq2_sub(s, s, bb); fq2_sub(s, s, cc); fq2_mul_by_xi(s, s); fq2_add(n0, s, aa);
    fq2_add(t, a.c0, a.c1); fq2_add(u, b.c0, b.c1); fq2_mul(s, t, u);
    fq2_sub(s, s, aa); fq2_sub(s, s, bb); fq2_mul_by_xi(t, cc); fq2_add(n1, s, t);
    fq2_add(t, a.c0, a.c2); fq2_add(u, b.c0, b.c2); fq2_mul(s, t, u);
    fq2_sub(s, s, aa); fq2_add(s, s, bb); fq2_sub(n2, s, cc);
    cpv(r.c0, n0); cpv(r.c1, n1); cpv(r.c2, n2);
}
BN_HD_NOINLINE void fq6_sqr(Fq6& r, const Fq6& a) {   // value of src/fields/fq6.rs:163-177
    Fq2 s0, s1, s2, s3, s4, t;
    fq2_sqr(s0, a.c0);
    fq2_mul(s1, a.c0, a.c1); fq2_dbl(s1, s1);
    fq2_sub(t, a.c0, a.c1); fq2_add(t, t, a.c2); fq2_sqr(s2, t);
    fq2_mul(s3, a.c1, a.c2); fq2_dbl(s3, s3);
    fq2_sqr(s4, a.c2);
    Fq2 n0, n1, n2;
    fq2_mul_by_xi(t, s3); fq2_add(n0, s0, t);
    fq2_mul_by_xi(t, s4); fq2_add(n1, s1, t);
    fq2_add(n2, s1, s2); fq2_add(n2, n2, s3); fq2_sub(n2, n2, s0); fq2_sub(n2, n2, s4);
    cpv(r.c0, n0); cpv(r.c1, n1); cpv(r.c2, n2);
}
BN_HD_NOINLINE void fq6_scale(Fq6& r, const Fq6& a, const Fq2& k) {   // src/fields/fq6.rs:117-123
    fq2_mul(r.c0, a.c0, k); fq2_mul(r.c1, a.c1, k); fq2_mul(r.c2, a.c2, k);
}
BN_HD_NOINLINE bool fq6_inv(Fq6& r, const Fq6& a) {   // src/fields/fq6.rs:179-191
    Fq2 c0, c1, c2, t, u;
    fq2_mul_by_xi(t, a.c2); fq2_mul(t, a.c1, t); fq2_sqr(c0, a.c0); fq2_sub(c0, c0, t);
    fq2_sqr(t, a.c2); fq2_mul_by_xi(t, t); fq2_mul(u, a.c0, a.c1); fq2_sub(c1, t, u);
    fq2_sqr(t, a.c1); fq2_mul(u, a.c0, a.c2); fq2_sub(c2, t, u);
    fq2_mul(t, a.c2, c1); fq2_mul(u, a.c1, c2); fq2_add(t, t, u); fq2_mul_by_xi(t, t);
    fq2_mul(u, a.c0, c0); fq2_add(t, t, u);
    if (!fq2_inv(u, t)) { r = fq6_zero(); return false; }
    fq2_mul(r.c0, u, c0); fq2_mul(r.c1, u, c1); fq2_mul(r.c2, u, c2);
    return true;
}
// Frobenius: conjugate each Fq2 coefficient for odd powers, multiply by xi^(i(q^k-1)/3); fq6.rs:125-131
BN_HD void fq6_frobenius(Fq6& r, const Fq6& a, int power) {
    const Fq2 c1tab[3] = {BN_C_FROB6_C1_1, BN_C_FROB6_C1_2, BN_C_FROB6_C1_3};
    const Fq2 c2tab[3] = {BN_C_FROB6_C2_1, BN_C_FROB6_C2_2, BN_C_FROB6_C2_3};
    Fq2 x0, x1, x2;
    if (power & 1) { fq2_conj(x0, a.c0); fq2_conj(x1, a.c1); fq2_conj(x2, a.c2); }
    else { cpv(x0, a.c0); cpv(x1, a.c1); cpv(x2, a.c2); }
    cpv(r.c0, x0);
    fq2_mul(r.c1, x1, c1tab[power - 1]);
    fq2_mul(r.c2, x2, c2tab[power - 1]);
}

// ---------------------------------------------------------------------------------------------
// Fq12 = Fq6[w]/(w^2 - v)
// ---------------------------------------------------------------------------------------------
BN_HD Fq12 fq12_one() { Fq12 r; r.c0 = fq6_one(); r.c1 = fq6_zero(); return r; }
BN_HD Fq12 fq12_zero() { Fq12 r; r.c0 = fq6_zero(); r.c1 = fq6_zero(); return r; }
BN_HD bool fq12_eq(const Fq12& a, const Fq12& b) {   // derived PartialEq on Gt, src/lib.rs:584
    const uint32_t* x = (const uint32_t*)&a;
    const uint32_t* y = (const uint32_t*)&b;
    uint32_t d = 0;
    for (int i = 0; i < 96; i++) d |= x[i] ^ y[i];
    return d == 0;
}
BN_HD_NOINLINE void fq12_mul(Fq12& r, const Fq12& a, const Fq12& b) {   // value of src/fields/fq12.rs:316-328
    Fq6 aa, bb, s, t, u;
    fq6_mul(aa, a.c0, b.c0);
    fq6_mul(bb, a.c1, b.c1);
    fq6_add(t, a.c0, a.c1);
    fq6_add(u, b.c0, b.c1);
    fq6_mul(s, t, u);
    fq6_sub(s, s, aa);
    fq6_sub(r.c1, s, bb);
    fq6_mul_by_v(t, bb);
    fq6_add(r.c0, t, aa);
}
BN_HD_NOINLINE void fq12_sqr(Fq12& r, const Fq12& a) {   // value of src/fields/fq12.rs:295-303
    Fq6 ab, t, u, s;
    fq6_mul(ab, a.c0, a.c1);
    fq6_mul_by_v(t, a.c1);
    fq6_add(t, t, a.c0);
    fq6_add(u, a.c0, a.c1);
    fq6_mul(s, t, u);
    fq6_sub(s, s, ab);
    fq6_mul_by_v(t, ab);
    fq6_sub(r.c0, s, t);
    fq6_dbl(r.c1, ab);
}
BN_HD void fq12_conj(Fq12& r, const Fq12& a) { if (&r != &a) cpv(r.c0, a.c0); fq6_neg(r.c1, a.c1); }   // unitary_inverse, fq12.rs:126-128
BN_HD_NOINLINE bool fq12_inv(Fq12& r, const Fq12& a) {   // src/fields/fq12.rs:305-313
    Fq6 t, u;
    fq6_sqr(t, a.c0);
    fq6_sqr(u, a.c1);
    fq6_mul_by_v(u, u);
    fq6_sub(t, t, u);
    if (!fq6_inv(u, t)) { r = fq12_zero(); return false; }
    fq6_mul(r.c0, a.c0, u);
    fq6_mul(t, a.c1, u);
    fq6_neg(r.c1, t);
    return true;
}
BN_HD_NOINLINE void fq12_frobenius(Fq12& r, const Fq12& a, int power) {   // src/fields/fq12.rs:112-119
    const Fq2 ctab[3] = {BN_C_FROB12_C1_1, BN_C_FROB12_C1_2, BN_C_FROB12_C1_3};
    Fq6 t;
    fq6_frobenius(r.c0, a.c0, power);
    fq6_frobenius(t, a.c1, power);
    fq6_scale(r.c1, t, ctab[power - 1]);
}

// f * (ell_0 + ell_vv v + ell_vw v w): sparse multiply; value of src/fields/fq12.rs:130-196.
// x0 = ell_0, x2 = ell_vv, x4 = ell_vw (already scaled by P.x / P.y by the caller).
BN_HD_NOINLINE void fq12_mul_by_024(Fq12& f, const Fq2& x0, const Fq2& x4, const Fq2& x2) {
    Fq2 z0, z1, z2, z3, z4, z5;
    cpv(z0, f.c0.c0); cpv(z1, f.c0.c1); cpv(z2, f.c0.c2); cpv(z3, f.c1.c0); cpv(z4, f.c1.c1); cpv(z5, f.c1.c2);
    Fq2 d0, d2, d4, s0, s1, t0, t1, t2, t3, t4;
    fq2_mul(d0, z0, x0);
    fq2_mul(d2, z2, x2);
    fq2_mul(d4, z4, x4);
    fq2_add(t2, z0, z4);
    fq2_add(t1, z0, z2);
    fq2_add(s0, z1, z3); fq2_add(s0, s0, z5);

    fq2_mul(s1, z1, x2);
    fq2_add(t3, s1, d4);
    fq2_mul_by_xi(t4, t3); fq2_add(f.c0.c0, t4, d0);          // z0'

    fq2_mul(t3, z5, x4);
    fq2_add(s1, s1, t3);
    fq2_add(t3, t3, d2);
    fq2_mul_by_xi(t4, t3);
    fq2_mul(t3, z1, x0);
    fq2_add(s1, s1, t3);
    fq2_add(f.c0.c1, t4, t3);                                  // z1'

    fq2_add(t0, x0, x2);
    fq2_mul(t3, t1, t0); fq2_sub(t3, t3, d0); fq2_sub(t3, t3, d2);
    fq2_mul(t4, z3, x4);
    fq2_add(s1, s1, t4);
    fq2_add(t0, z2, z4);
    fq2_add(f.c0.c2, t3, t4);                                  // z2'

    fq2_add(t1, x2, x4);
    fq2_mul(t3, t0, t1); fq2_sub(t3, t3, d2); fq2_sub(t3, t3, d4);
    fq2_mul_by_xi(t4, t3);
    fq2_mul(t3, z3, x0);
    fq2_add(s1, s1, t3);
    fq2_add(f.c1.c0, t4, t3);                                  // z3'

    fq2_mul(t3, z5, x2);
    fq2_add(s1, s1, t3);
    fq2_mul_by_xi(t4, t3);
    fq2_add(t0, x0, x4);
    fq2_mul(t3, t2, t0); fq2_sub(t3, t3, d0); fq2_sub(t3, t3, d4);
    fq2_add(f.c1.c1, t4, t3);                                  // z4'

    fq2_add(t0, x0, x2); fq2_add(t0, t0, x4);
    fq2_mul(t3, s0, t0);
    fq2_sub(f.c1.c2, t3, s1);                                  // z5'
}

// Granger-Scott squaring in the cyclotomic subgroup; value of src/fields/fq12.rs:198-247.
// One "Fq4 squaring" (a, b) -> (a^2 + xi b^2, 2ab) per coefficient pair.
BN_HD void fq4_sqr(Fq2& r0, Fq2& r1, const Fq2& a, const Fq2& b) {
    Fq2 tmp, s, t, u;
    fq2_mul(tmp, a, b);
    fq2_add(s, a, b);
    fq2_mul_by_xi(t, b); fq2_add(t, t, a);
    fq2_mul(s, s, t);
    fq2_sub(s, s, tmp);
    fq2_mul_by_xi(u, tmp);
    fq2_sub(r0, s, u);
    fq2_dbl(r1, tmp);
}
BN_HD_NOINLINE void fq12_cyclotomic_sqr(Fq12& r, const Fq12& a) {
    Fq2 z0, z1, z2, z3, z4, z5;
    cpv(z0, a.c0.c0); cpv(z4, a.c0.c1); cpv(z3, a.c0.c2); cpv(z2, a.c1.c0); cpv(z1, a.c1.c1); cpv(z5, a.c1.c2);
    Fq2 t0, t1, t2, t3, t4, t5, tmp;
    fq4_sqr(t0, t1, z0, z1);
    fq4_sqr(t2, t3, z2, z3);
    fq4_sqr(t4, t5, z4, z5);
    fq2_sub(z0, t0, z0); fq2_dbl(z0, z0); fq2_add(z0, z0, t0);
    fq2_add(z1, t1, z1); fq2_dbl(z1, z1); fq2_add(z1, z1, t1);
    fq2_mul_by_xi(tmp, t5);
    fq2_add(z2, tmp, z2); fq2_dbl(z2, z2); fq2_add(z2, z2, tmp);
    fq2_sub(z3, t4, z3); fq2_dbl(z3, z3); fq2_add(z3, z3, t4);
    fq2_sub(z4, t2, z4); fq2_dbl(z4, z4); fq2_add(z4, z4, t2);
    fq2_add(z5, t3, z5); fq2_dbl(z5, z5); fq2_add(z5, z5, t3);
    cpv(r.c0.c0, z0); cpv(r.c0.c1, z4); cpv(r.c0.c2, z3);
    cpv(r.c1.c0, z2); cpv(r.c1.c1, z1); cpv(r.c1.c2, z5);
}
// conj(a^u), u = 4965661367192848881 (63 bits); value of src/fields/fq12.rs:121-124 + :249-266.
// The reference runs plain square-and-multiply (62 squarings, 28 multiplications incl. one by 1).  a lies in
// the cyclotomic subgroup here (a^-1 = conj(a) is free), so the same power is reached with the width-3 signed
// window recoding u = sum d_i 2^i, d_i in {0, +-1, +-3}: 62 squarings, 17 multiplications + 2 for a^3.
// Bit i of NZ: d_i != 0; of NEG: d_i < 0; of M3: |d_i| = 3 (generated and checked against u by tools/gen_constants.py).
#define BN_U_WNAF3_NZ 0x4908924444891211ull
#define BN_U_WNAF3_NEG 0x0108000400880210ull
#define BN_U_WNAF3_M3 0x0108804404880200ull
// Non-aliasing variants for the hot loops (r must not overlap the inputs): the output doubles as scratch, so
// the frame is 2 Fq6 instead of 5 and no coefficient is copied.  conj_b multiplies by conj(b) = b^-1 for free.
BN_HD_NOINLINE void fq12_mul_na(Fq12& r, const Fq12& a, const Fq12& b, bool conj_b) {
    Fq6 aa, bb;
    fq6_add(r.c1, a.c0, a.c1);
    if (conj_b) fq6_sub(r.c0, b.c0, b.c1); else fq6_add(r.c0, b.c0, b.c1);
    fq6_mul(r.c1, r.c1, r.c0);                     // (a0 + a1)(b0 +- b1)
    fq6_mul(aa, a.c0, b.c0);
    fq6_mul(bb, a.c1, b.c1);
    fq6_sub(r.c1, r.c1, aa);
    fq6_mul_by_v(r.c0, bb);
    if (conj_b) { fq6_add(r.c1, r.c1, bb); fq6_sub(r.c0, aa, r.c0); }
    else { fq6_sub(r.c1, r.c1, bb); fq6_add(r.c0, r.c0, aa); }
}
BN_HD_NOINLINE void fq12_cyclotomic_sqr_na(Fq12& r, const Fq12& a) {
    // reference labels (fq12.rs:199-204): z0 = c0.c0, z4 = c0.c1, z3 = c0.c2, z2 = c1.c0, z1 = c1.c1, z5 = c1.c2.
    // One coefficient pair at a time, its two results written before the next pair starts: two live temporaries
    // instead of six, and each input coefficient is re-read while it is still in L1.
    Fq2 t0, t1;
    fq4_sqr(t0, t1, a.c0.c0, a.c1.c1);
    fq2_sub(r.c0.c0, t0, a.c0.c0); fq2_dbl(r.c0.c0, r.c0.c0); fq2_add(r.c0.c0, r.c0.c0, t0);   // z0' = 3 t0 - 2 z0
    fq2_add(r.c1.c1, t1, a.c1.c1); fq2_dbl(r.c1.c1, r.c1.c1); fq2_add(r.c1.c1, r.c1.c1, t1);   // z1' = 3 t1 + 2 z1
    fq4_sqr(t0, t1, a.c1.c0, a.c0.c2);                                                         // (t2, t3)
    fq2_sub(r.c0.c1, t0, a.c0.c1); fq2_dbl(r.c0.c1, r.c0.c1); fq2_add(r.c0.c1, r.c0.c1, t0);   // z4' = 3 t2 - 2 z4
    fq2_add(r.c1.c2, t1, a.c1.c2); fq2_dbl(r.c1.c2, r.c1.c2); fq2_add(r.c1.c2, r.c1.c2, t1);   // z5' = 3 t3 + 2 z5
    fq4_sqr(t0, t1, a.c0.c1, a.c1.c2);                                                         // (t4, t5)
    fq2_sub(r.c0.c2, t0, a.c0.c2); fq2_dbl(r.c0.c2, r.c0.c2); fq2_add(r.c0.c2, r.c0.c2, t0);   // z3' = 3 t4 - 2 z3
    fq2_mul_by_xi(t1, t1);
    fq2_add(r.c1.c0, t1, a.c1.c0); fq2_dbl(r.c1.c0, r.c1.c0); fq2_add(r.c1.c0, r.c1.c0, t1);   // z2' = 3 xi t5 + 2 z2
}
// Width-4 recoding of the same u: digits in {0, +-1, +-3, +-5, +-7}, 14 non-zero (13 multiplications + 3 for
// a^3, a^5, a^7).  Bits W4_I0/W4_I1 of position i give (|d_i| - 1) / 2.  The default; -DBN_U_WINDOW=3 selects the width-3 digits.
#define BN_U_WNAF4_NZ 0x4108844442110211ull
#define BN_U_WNAF4_NEG 0x0008004400010010ull
#define BN_U_WNAF4_I0 0x0008800400110000ull
#define BN_U_WNAF4_I1 0x0100044002110200ull
#ifndef BN_U_WINDOW
#define BN_U_WINDOW 4
#endif
BN_HD_NOINLINE void fq12_exp_by_neg_z(Fq12& r, const Fq12& a) {
    // Two accumulators used alternately (every step reads one and writes the other).  The buffers are named
    // statically in both branches -- a pointer table here makes nvcc lose the local address space of EVERY
    // Fq2-level body (generic LD/ST instead of LDL/STL: -5 % on all kernels, measured).
#if BN_U_WINDOW == 4
    Fq12 a3, a5, a7, buf0, buf1;
    fq12_cyclotomic_sqr_na(buf0, a);               // a^2
    fq12_mul_na(a3, buf0, a, false);
    fq12_mul_na(a5, buf0, a3, false);
    fq12_mul_na(a7, buf0, a5, false);
    bool in0 = true;                               // top digit (bit 62) is +1; buf0 already holds a^2 (step i = 61)
    for (int i = 61; i >= 0; i--) {
        if (i != 61) {
            if (in0) fq12_cyclotomic_sqr_na(buf1, buf0); else fq12_cyclotomic_sqr_na(buf0, buf1);
            in0 = !in0;
        }
        if ((BN_U_WNAF4_NZ >> i) & 1) {
            bool neg = ((BN_U_WNAF4_NEG >> i) & 1) != 0;
            int idx = (int)((BN_U_WNAF4_I0 >> i) & 1) | ((int)((BN_U_WNAF4_I1 >> i) & 1) << 1);
            if (idx == 0) { if (in0) fq12_mul_na(buf1, buf0, a, neg); else fq12_mul_na(buf0, buf1, a, neg); }
            else if (idx == 1) { if (in0) fq12_mul_na(buf1, buf0, a3, neg); else fq12_mul_na(buf0, buf1, a3, neg); }
            else if (idx == 2) { if (in0) fq12_mul_na(buf1, buf0, a5, neg); else fq12_mul_na(buf0, buf1, a5, neg); }
            else { if (in0) fq12_mul_na(buf1, buf0, a7, neg); else fq12_mul_na(buf0, buf1, a7, neg); }
            in0 = !in0;
        }
    }
    if (in0) fq12_conj(r, buf0); else fq12_conj(r, buf1);
#else
    Fq12 a3, buf0, buf1;
    fq12_cyclotomic_sqr_na(buf0, a);
    fq12_mul_na(a3, buf0, a, false);
    fq12_cyclotomic_sqr_na(buf0, a);               // top digit (bit 62) is +1; step i = 61 squares a itself
    bool in0 = true;                               // the current value lives in buf0
    for (int i = 61; i >= 0; i--) {
        if (i != 61) {
            if (in0) fq12_cyclotomic_sqr_na(buf1, buf0); else fq12_cyclotomic_sqr_na(buf0, buf1);
            in0 = !in0;
        }
        if ((BN_U_WNAF3_NZ >> i) & 1) {
            bool neg = ((BN_U_WNAF3_NEG >> i) & 1) != 0;
            if ((BN_U_WNAF3_M3 >> i) & 1) {
                if (in0) fq12_mul_na(buf1, buf0, a3, neg); else fq12_mul_na(buf0, buf1, a3, neg);
            } else {
                if (in0) fq12_mul_na(buf1, buf0, a, neg); else fq12_mul_na(buf0, buf1, a, neg);
            }
            in0 = !in0;
        }
    }
    if (in0) fq12_conj(r, buf0); else fq12_conj(r, buf1);
#endif
}
// easy part f^((q^6-1)(q^2+1)); false iff f == 0.  src/fields/fq12.rs:62-73
BN_HD_NOINLINE bool fq12_final_exp_first_chunk(Fq12& r, const Fq12& f) {
    Fq12 b, a, c;
    if (!fq12_inv(b, f)) { r = fq12_zero(); return false; }
    fq12_conj(a, f);
    fq12_mul(c, a, b);
    fq12_frobenius(a, c, 2);
    fq12_mul(r, a, c);
    return true;
}
// hard part, the reference's addition chain; src/fields/fq12.rs:75-105 (same step names)
BN_HD_NOINLINE void fq12_final_exp_last_chunk(Fq12& out, const Fq12& x) {
    Fq12 a, b, c, d, e, f, g, k, l, t;
    fq12_exp_by_neg_z(a, x);
    fq12_cyclotomic_sqr(b, a);
    fq12_cyclotomic_sqr(c, b);
    fq12_mul(d, c, b);
    fq12_exp_by_neg_z(e, d);
    fq12_cyclotomic_sqr(f, e);
    fq12_exp_by_neg_z(g, f);
    fq12_conj(d, d);              // h
    fq12_conj(g, g);              // i
    fq12_mul(t, g, e);            // j
    fq12_mul(k, t, d);            // k
    fq12_mul(l, k, b);            // l
    fq12_mul(t, k, e);            // m
    fq12_mul(t, x, t);            // n
    fq12_frobenius(a, l, 1);      // o
    fq12_mul(t, a, t);            // p
    fq12_frobenius(a, k, 2);      // q
    fq12_mul(t, a, t);            // r
    fq12_conj(a, x);              // s
    fq12_mul(a, a, l);            // t
    fq12_frobenius(b, a, 3);      // u
    fq12_mul(out, b, t);          // v
}
BN_HD bool fq12_final_exponentiation(Fq12& r, const Fq12& f) {   // src/fields/fq12.rs:107-110
    Fq12 t;
    if (!fq12_final_exp_first_chunk(t, f)) { r = fq12_zero(); return false; }
    fq12_final_exp_last_chunk(r, t);
    return true;
}
// FieldElement::pow over the 256 bits of a canonical scalar (Gt::pow, src/lib.rs:592 -> fields/mod.rs:35-46).
// Leading zero bits only square 1, so skipping them gives the same value.
BN_HD_NOINLINE void fq12_pow(Fq12& r, const Fq12& a, const uint32_t (&k)[8]) {
    Fq12 acc = fq12_one();
    bool started = false;
    for (int i = 255; i >= 0; i--) {
        if (started) fq12_sqr(acc, acc);
        if ((k[i >> 5] >> (i & 31)) & 1) {
            if (started) fq12_mul(acc, acc, a); else { acc = a; started = true; }
        }
    }
    r = acc;
}

}  // namespace bn
