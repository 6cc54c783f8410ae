// BN254 base field Fq on 8 x 32-bit limbs, Montgomery form (R = 2^256), always fully reduced.
//
// Replaces, on the device, the reference's U256 arithmetic (src/arith.rs:281-316 add/sub/neg,
// :300-306 + :525-545 mul/mul_reduce) and the Fq newtype (src/fields/fp.rs:7-222).  The reference
// uses 2 x u128 limbs and a radix-2^128 Montgomery step; both compute a*b*2^-256 mod q with a final
// conditional subtraction, so the 32-byte little-endian memory image of every result is identical.
//
// Two back ends behind one interface:
//   * device (__CUDA_ARCH__): the generated single-block PTX carry chains in fq_ptx.cuh
//     (mad.lo.cc / madc.hi.cc -> IMAD.WIDE.U32[.X] on sm_100a);
//   * host (BN_HOST_SIM, tests only): portable 64-bit C++ with the same results.  The shipped
//     library never executes the host back end: every exported entry point launches kernels.
#pragma once
#include <stdint.h>

#if defined(__CUDACC__)
#define BN_HD __host__ __device__ __forceinline__
#define BN_HD_NOINLINE __host__ __device__ __noinline__
#include "fq_ptx.cuh"
#else
#define BN_HD inline
#define BN_HD_NOINLINE
#endif

namespace bn {

struct alignas(16) Fq {
    uint32_t v[8];
};
struct alignas(16) FqWide {  // 512-bit unreduced product (lazy reduction)
    uint32_t v[16];
};

// q, little-endian 32-bit limbs (src/fields/fp.rs:197-202)
#define BN_Q0 0xd87cfd47u
#define BN_Q1 0x3c208c16u
#define BN_Q2 0x6871ca8du
#define BN_Q3 0x97816a91u
#define BN_Q4 0x8181585du
#define BN_Q5 0xb85045b6u
#define BN_Q6 0xe131a029u
#define BN_Q7 0x30644e72u
#define BN_QINV32 0xe4866389u  // -q^-1 mod 2^32 (low word of src/fields/fp.rs:221)

BN_HD uint32_t fq_modulus_limb(int i) {
    switch (i) {
        case 0: return BN_Q0; case 1: return BN_Q1; case 2: return BN_Q2; case 3: return BN_Q3;
        case 4: return BN_Q4; case 5: return BN_Q5; case 6: return BN_Q6; default: return BN_Q7;
    }
}

BN_HD bool fq_is_zero(const Fq& a) {
    uint32_t x = 0;
#pragma unroll
    for (int i = 0; i < 8; i++) x |= a.v[i];
    return x == 0;
}
BN_HD bool fq_eq(const Fq& a, const Fq& b) {
    uint32_t x = 0;
#pragma unroll
    for (int i = 0; i < 8; i++) x |= a.v[i] ^ b.v[i];
    return x == 0;
}
BN_HD void fq_set_zero(Fq& a) {
#pragma unroll
    for (int i = 0; i < 8; i++) a.v[i] = 0;
}

// 128-bit loads/stores of field elements that live in memory (local-memory Fq12 state, global I/O):
// the out-of-line Fq2-level bodies read their operands with 2 x LD.128 per Fq instead of 8 scalar loads.
BN_HD Fq fq_ld(const Fq& p) {
#if defined(__CUDA_ARCH__)
    const uint4* q = reinterpret_cast<const uint4*>(&p);
    uint4 a = q[0], b = q[1];
    Fq r;
    r.v[0] = a.x; r.v[1] = a.y; r.v[2] = a.z; r.v[3] = a.w; r.v[4] = b.x; r.v[5] = b.y; r.v[6] = b.z; r.v[7] = b.w;
    return r;
#else
    return p;
#endif
}
BN_HD void fq_st(Fq& p, const Fq& v) {
#if defined(__CUDA_ARCH__)
    uint4* q = reinterpret_cast<uint4*>(&p);
    q[0] = make_uint4(v.v[0], v.v[1], v.v[2], v.v[3]);
    q[1] = make_uint4(v.v[4], v.v[5], v.v[6], v.v[7]);
#else
    p = v;
#endif
}

#if !defined(__CUDA_ARCH__)
// ---------------------------------------------------------------------------------------------
// host back end (test-only simulation of the device arithmetic)
// ---------------------------------------------------------------------------------------------
namespace hostimpl {
inline bool geq_q(const uint32_t* a) {
    for (int i = 7; i >= 0; i--) {
        uint32_t q = fq_modulus_limb(i);
        if (a[i] != q) return a[i] > q;
    }
    return true;
}
inline void sub_q(uint32_t* a) {
    int64_t br = 0;
    for (int i = 0; i < 8; i++) {
        int64_t d = (int64_t)a[i] - fq_modulus_limb(i) + br;
        a[i] = (uint32_t)d;
        br = d >> 32;
    }
}
inline void mulwide(uint32_t* t, const uint32_t* a, const uint32_t* b) {
    for (int i = 0; i < 16; i++) t[i] = 0;
    for (int i = 0; i < 8; i++) {
        uint64_t c = 0;
        for (int j = 0; j < 8; j++) {
            uint64_t s = (uint64_t)a[j] * b[i] + t[i + j] + c;
            t[i + j] = (uint32_t)s;
            c = s >> 32;
        }
        t[i + 8] = (uint32_t)c;
    }
}
inline void redc(uint32_t* r, const uint32_t* tin) {
    uint32_t t[17];
    for (int i = 0; i < 16; i++) t[i] = tin[i];
    t[16] = 0;
    for (int i = 0; i < 8; i++) {
        uint32_t k = t[i] * BN_QINV32;
        uint64_t c = 0;
        for (int j = 0; j < 8; j++) {
            uint64_t s = (uint64_t)k * fq_modulus_limb(j) + t[i + j] + c;
            t[i + j] = (uint32_t)s;
            c = s >> 32;
        }
        for (int j = i + 8; c && j < 17; j++) {
            uint64_t s = (uint64_t)t[j] + c;
            t[j] = (uint32_t)s;
            c = s >> 32;
        }
    }
    for (int i = 0; i < 8; i++) r[i] = t[8 + i];
    if (t[16] || geq_q(r)) sub_q(r);
}
}  // namespace hostimpl
#endif

BN_HD void fq_add(Fq& r, const Fq& a, const Fq& b) {  // src/arith.rs:281-287
#if defined(__CUDA_ARCH__)
    uint32_t t[8];
    fq_add_ptx(t, a.v, b.v);
#pragma unroll
    for (int i = 0; i < 8; i++) r.v[i] = t[i];
#else
    uint32_t t[8];
    uint64_t c = 0;
    for (int i = 0; i < 8; i++) {
        uint64_t s = (uint64_t)a.v[i] + b.v[i] + c;
        t[i] = (uint32_t)s;
        c = s >> 32;
    }
    if (hostimpl::geq_q(t)) hostimpl::sub_q(t);
    for (int i = 0; i < 8; i++) r.v[i] = t[i];
#endif
}
BN_HD void fq_sub(Fq& r, const Fq& a, const Fq& b) {  // src/arith.rs:290-296
#if defined(__CUDA_ARCH__)
    uint32_t t[8];
    fq_sub_ptx(t, a.v, b.v);
#pragma unroll
    for (int i = 0; i < 8; i++) r.v[i] = t[i];
#else
    uint32_t t[8];
    int64_t br = 0;
    for (int i = 0; i < 8; i++) {
        int64_t d = (int64_t)a.v[i] - b.v[i] + br;
        t[i] = (uint32_t)d;
        br = d >> 32;
    }
    if (br) {
        uint64_t c = 0;
        for (int i = 0; i < 8; i++) {
            uint64_t s = (uint64_t)t[i] + fq_modulus_limb(i) + c;
            t[i] = (uint32_t)s;
            c = s >> 32;
        }
    }
    for (int i = 0; i < 8; i++) r.v[i] = t[i];
#endif
}
// r = 9a + b  /  9a - b  mod q in one pass (the halves of a multiplication by xi = 9 + u, src/fields/fq2.rs:55-57):
// t = 9a + b' < 10q over 9 limbs (b' = b or q - b), quotient estimate k from the top 16 bits with
// k <= floor(t/q) <= k + 1, r = t - k q (< 2q, formed modulo 2^256), one conditional subtraction.
// tools/gen_fq_ptx.py sweeps the estimate exhaustively over its 16-bit input.
template <bool MINUS> BN_HD void fq_lin9(Fq& r, const Fq& a, const Fq& b) {
#if defined(__CUDA_ARCH__)
    uint32_t t[8];
    if (MINUS) fq_9a_minus_b_ptx(t, a.v, b.v); else fq_9a_plus_b_ptx(t, a.v, b.v);
#pragma unroll
    for (int i = 0; i < 8; i++) r.v[i] = t[i];
#else
    uint32_t bp[8], t[9];
    if (MINUS) {
        int64_t br = 0;
        for (int i = 0; i < 8; i++) {
            int64_t d = (int64_t)fq_modulus_limb(i) - b.v[i] + br;
            bp[i] = (uint32_t)d;
            br = d >> 32;
        }
    } else {
        for (int i = 0; i < 8; i++) bp[i] = b.v[i];
    }
    uint64_t c = 0;
    for (int i = 0; i < 8; i++) {
        uint64_t s = (uint64_t)a.v[i] * 9u + bp[i] + c;
        t[i] = (uint32_t)s;
        c = s >> 32;
    }
    t[8] = (uint32_t)c;
    uint32_t th = (t[8] << 14) | (t[7] >> 18);
    uint32_t k = (uint32_t)(((uint64_t)th * 1386367u) >> 32);     // floor(2^32 / 3098), 3097 = q >> 242
    int64_t br = 0;
    uint64_t mc = 0;
    uint32_t o[8];
    for (int i = 0; i < 8; i++) {
        uint64_t m = (uint64_t)fq_modulus_limb(i) * k + mc;
        mc = m >> 32;
        int64_t d = (int64_t)t[i] - (int64_t)(uint32_t)m + br;
        o[i] = (uint32_t)d;
        br = d >> 32;
    }
    if (hostimpl::geq_q(o)) hostimpl::sub_q(o);
    for (int i = 0; i < 8; i++) r.v[i] = o[i];
#endif
}
// ---- sums without a reduction per term (the quad-lane executor, bn_coop4.cuh) --------------------------------------
// A 288-bit two's-complement accumulator: up to 8 signed terms in [0, q) and one multiplication by 9 stay far inside it.
// acc9_add: a += x (mask = 0) or a -= x (mask = 0xffffffff) as ~x + 1.  acc9_reduce: (a + offset_q * q) mod q for a total in
// [0, 160 q) (fq_red9_ptx: quotient estimate from the top bits + one conditional subtraction, tools/gen_fq_ptx.py).
struct Acc9 { uint32_t v[9]; };
BN_HD void acc9_zero(Acc9& a) {
#pragma unroll
    for (int i = 0; i < 9; i++) a.v[i] = 0;
}
BN_HD void acc9_add(Acc9& a, const Fq& x, uint32_t mask) {
#if defined(__CUDA_ARCH__)
    asm("{\n\t.reg .u32 d;\n\t"
        "add.cc.u32 d, %17, 1;\n\t"                  // carry in = 1 when subtracting
        "addc.cc.u32 %0, %0, %9;\n\taddc.cc.u32 %1, %1, %10;\n\taddc.cc.u32 %2, %2, %11;\n\taddc.cc.u32 %3, %3, %12;\n\t"
        "addc.cc.u32 %4, %4, %13;\n\taddc.cc.u32 %5, %5, %14;\n\taddc.cc.u32 %6, %6, %15;\n\taddc.cc.u32 %7, %7, %16;\n\t"
        "addc.u32 %8, %8, %17;\n\t}"
        : "+r"(a.v[0]), "+r"(a.v[1]), "+r"(a.v[2]), "+r"(a.v[3]), "+r"(a.v[4]), "+r"(a.v[5]), "+r"(a.v[6]), "+r"(a.v[7]), "+r"(a.v[8])
        : "r"(x.v[0] ^ mask), "r"(x.v[1] ^ mask), "r"(x.v[2] ^ mask), "r"(x.v[3] ^ mask), "r"(x.v[4] ^ mask), "r"(x.v[5] ^ mask),
          "r"(x.v[6] ^ mask), "r"(x.v[7] ^ mask), "r"(mask));
#else
    uint64_t c = mask & 1u;
    for (int i = 0; i < 9; i++) {
        uint64_t s = (uint64_t)a.v[i] + (i < 8 ? (x.v[i] ^ mask) : mask) + c;
        a.v[i] = (uint32_t)s;
        c = s >> 32;
    }
#endif
}
// a += b (mod 2^288)
BN_HD void acc9_addv(Acc9& a, const Acc9& b) {
#if defined(__CUDA_ARCH__)
    asm("add.cc.u32 %0, %0, %9;\n\taddc.cc.u32 %1, %1, %10;\n\taddc.cc.u32 %2, %2, %11;\n\taddc.cc.u32 %3, %3, %12;\n\t"
        "addc.cc.u32 %4, %4, %13;\n\taddc.cc.u32 %5, %5, %14;\n\taddc.cc.u32 %6, %6, %15;\n\taddc.cc.u32 %7, %7, %16;\n\t"
        "addc.u32 %8, %8, %17;"
        : "+r"(a.v[0]), "+r"(a.v[1]), "+r"(a.v[2]), "+r"(a.v[3]), "+r"(a.v[4]), "+r"(a.v[5]), "+r"(a.v[6]), "+r"(a.v[7]), "+r"(a.v[8])
        : "r"(b.v[0]), "r"(b.v[1]), "r"(b.v[2]), "r"(b.v[3]), "r"(b.v[4]), "r"(b.v[5]), "r"(b.v[6]), "r"(b.v[7]), "r"(b.v[8]));
#else
    uint64_t c = 0;
    for (int i = 0; i < 9; i++) {
        uint64_t s = (uint64_t)a.v[i] + b.v[i] + c;
        a.v[i] = (uint32_t)s;
        c = s >> 32;
    }
#endif
}
// a += 9 b (mod 2^288): b + (b << 3)
BN_HD void acc9_add9(Acc9& a, const Acc9& b) {
    Acc9 s;
    s.v[0] = b.v[0] << 3;
#pragma unroll
    for (int i = 1; i < 9; i++) s.v[i] = (b.v[i] << 3) | (b.v[i - 1] >> 29);
    acc9_addv(a, b);
    acc9_addv(a, s);
}
// a += x, a -= x for a canonical x (fixed sign: no masking), a = b << k (mod 2^288)
BN_HD void acc9_addu(Acc9& a, const Fq& x) {
#if defined(__CUDA_ARCH__)
    asm("add.cc.u32 %0, %0, %9;\n\taddc.cc.u32 %1, %1, %10;\n\taddc.cc.u32 %2, %2, %11;\n\taddc.cc.u32 %3, %3, %12;\n\t"
        "addc.cc.u32 %4, %4, %13;\n\taddc.cc.u32 %5, %5, %14;\n\taddc.cc.u32 %6, %6, %15;\n\taddc.cc.u32 %7, %7, %16;\n\t"
        "addc.u32 %8, %8, 0;"
        : "+r"(a.v[0]), "+r"(a.v[1]), "+r"(a.v[2]), "+r"(a.v[3]), "+r"(a.v[4]), "+r"(a.v[5]), "+r"(a.v[6]), "+r"(a.v[7]), "+r"(a.v[8])
        : "r"(x.v[0]), "r"(x.v[1]), "r"(x.v[2]), "r"(x.v[3]), "r"(x.v[4]), "r"(x.v[5]), "r"(x.v[6]), "r"(x.v[7]));
#else
    acc9_add(a, x, 0u);
#endif
}
BN_HD void acc9_subu(Acc9& a, const Fq& x) {
#if defined(__CUDA_ARCH__)
    asm("sub.cc.u32 %0, %0, %9;\n\tsubc.cc.u32 %1, %1, %10;\n\tsubc.cc.u32 %2, %2, %11;\n\tsubc.cc.u32 %3, %3, %12;\n\t"
        "subc.cc.u32 %4, %4, %13;\n\tsubc.cc.u32 %5, %5, %14;\n\tsubc.cc.u32 %6, %6, %15;\n\tsubc.cc.u32 %7, %7, %16;\n\t"
        "subc.u32 %8, %8, 0;"
        : "+r"(a.v[0]), "+r"(a.v[1]), "+r"(a.v[2]), "+r"(a.v[3]), "+r"(a.v[4]), "+r"(a.v[5]), "+r"(a.v[6]), "+r"(a.v[7]), "+r"(a.v[8])
        : "r"(x.v[0]), "r"(x.v[1]), "r"(x.v[2]), "r"(x.v[3]), "r"(x.v[4]), "r"(x.v[5]), "r"(x.v[6]), "r"(x.v[7]));
#else
    acc9_add(a, x, 0xffffffffu);
#endif
}
BN_HD void acc9_subv(Acc9& a, const Acc9& b) {
#if defined(__CUDA_ARCH__)
    asm("sub.cc.u32 %0, %0, %9;\n\tsubc.cc.u32 %1, %1, %10;\n\tsubc.cc.u32 %2, %2, %11;\n\tsubc.cc.u32 %3, %3, %12;\n\t"
        "subc.cc.u32 %4, %4, %13;\n\tsubc.cc.u32 %5, %5, %14;\n\tsubc.cc.u32 %6, %6, %15;\n\tsubc.cc.u32 %7, %7, %16;\n\t"
        "subc.u32 %8, %8, %17;"
        : "+r"(a.v[0]), "+r"(a.v[1]), "+r"(a.v[2]), "+r"(a.v[3]), "+r"(a.v[4]), "+r"(a.v[5]), "+r"(a.v[6]), "+r"(a.v[7]), "+r"(a.v[8])
        : "r"(b.v[0]), "r"(b.v[1]), "r"(b.v[2]), "r"(b.v[3]), "r"(b.v[4]), "r"(b.v[5]), "r"(b.v[6]), "r"(b.v[7]), "r"(b.v[8]));
#else
    long long br = 0;
    for (int i = 0; i < 9; i++) {
        long long d = (long long)a.v[i] - (long long)b.v[i] + br;
        a.v[i] = (uint32_t)d;
        br = d >> 32;
    }
#endif
}
template <int K> BN_HD void acc9_shl(Acc9& r, const Acc9& a) {
    uint32_t prev = 0;
#pragma unroll
    for (int i = 0; i < 9; i++) { const uint32_t cur = a.v[i]; r.v[i] = (cur << K) | (prev >> (32 - K)); prev = cur; }
}
BN_HD void acc9_set(Acc9& a, const Fq& x) {
#pragma unroll
    for (int i = 0; i < 8; i++) a.v[i] = x.v[i];
    a.v[8] = 0;
}
// limb i of m q for m = 8 and m = 80 (the offsets that make the sums of the MUL and LIN steps non-negative)
BN_HD uint32_t acc9_offset_limb(int m80, int i) {
    // 8 q and 80 q over nine limbs
    const uint32_t q8[9] = {0xc3e7ea38u, 0xe10460b6u, 0x438e5469u, 0xbc0b548bu, 0x0c0ac2ecu, 0xc2822db4u, 0x098d014du, 0x83227397u, 0x00000001u};
    const uint32_t q80[9] = {0xa70f2630u, 0xca2bc723u, 0xa38f4c22u, 0x58714d70u, 0x786b9d3fu, 0x9915c908u, 0x5f820d09u, 0x1f5883e6u, 0x0000000fu};
    return m80 ? q80[i] : q8[i];
}
template <bool M80> BN_HD void acc9_reduce(Fq& r, const Acc9& a) {
    Acc9 t = a, off;
#pragma unroll
    for (int i = 0; i < 9; i++) off.v[i] = acc9_offset_limb(M80 ? 1 : 0, i);
    acc9_addv(t, off);
#if defined(__CUDA_ARCH__)
    uint32_t o[8];
    fq_red9_ptx(o, t.v);
#pragma unroll
    for (int i = 0; i < 8; i++) r.v[i] = o[i];
#else
    uint32_t th = (t.v[8] << 14) | (t.v[7] >> 18);
    uint32_t k = (uint32_t)(((uint64_t)th * 1386367u) >> 32);
    int64_t br = 0;
    uint64_t mc = 0;
    uint32_t o[8];
    for (int i = 0; i < 8; i++) {
        uint64_t m = (uint64_t)fq_modulus_limb(i) * k + mc;
        mc = m >> 32;
        int64_t d = (int64_t)t.v[i] - (int64_t)(uint32_t)m + br;
        o[i] = (uint32_t)d;
        br = d >> 32;
    }
    if (hostimpl::geq_q(o)) hostimpl::sub_q(o);
    for (int i = 0; i < 8; i++) r.v[i] = o[i];
#endif
}

// out_c component = 3 (S - 10 m +- m') - 2 z as ONE lazily accumulated 288-bit sum, reduced once (|sum| <= 38 q): 129 instructions
// against 162 for the chain lin9, sub, sub, sub, dbl, add with a conditional correction each (hard part 9.61 -> 9.56 ms)
BN_HD void cyc_outc(Fq& out, const Fq& s, const Fq& m, const Fq& mx, bool plus, const Fq& z) {
    Acc9 t, u, v;
    acc9_set(t, s);
    if (plus) acc9_addu(t, mx); else acc9_subu(t, mx);
    acc9_set(u, m);
    acc9_shl<3>(v, u); acc9_subv(t, v);
    acc9_shl<1>(v, u); acc9_subv(t, v);                      // t = S +- m' - 10 m
    acc9_shl<1>(v, t); acc9_addv(t, v);                      // 3 t
    acc9_set(u, z);
    acc9_shl<1>(v, u); acc9_subv(t, v);                      // - 2 z
    acc9_reduce<true>(out, t);
}
BN_HD void fq_neg(Fq& r, const Fq& a) {  // src/arith.rs:309-316 (0 -> 0)
    Fq z;
    fq_set_zero(z);
    fq_sub(r, z, a);
}
BN_HD void fq_dbl(Fq& r, const Fq& a) { fq_add(r, a, a); }

BN_HD void fq_mul(Fq& r, const Fq& a, const Fq& b) {  // src/arith.rs:300-306, :525-545
#if defined(__CUDA_ARCH__)
    uint32_t t[8];
    fq_mul_ptx(t, a.v, b.v);
#pragma unroll
    for (int i = 0; i < 8; i++) r.v[i] = t[i];
#else
    uint32_t w[16], t[8];
    hostimpl::mulwide(w, a.v, b.v);
    hostimpl::redc(t, w);
    for (int i = 0; i < 8; i++) r.v[i] = t[i];
#endif
}
BN_HD void fq_sqr(Fq& r, const Fq& a) { fq_mul(r, a, a); }

// ---- lazy-reduction helpers: 512-bit products, reduced once per output coefficient ----------
BN_HD void fq_mulwide(FqWide& t, const Fq& a, const Fq& b) {
#if defined(__CUDA_ARCH__)
    fq_mulwide_ptx(t.v, a.v, b.v);
#else
    hostimpl::mulwide(t.v, a.v, b.v);
#endif
}
BN_HD void fq_redc(Fq& r, const FqWide& t) {  // t < q*2^256
#if defined(__CUDA_ARCH__)
    uint32_t o[8];
    fq_redc_ptx(o, t.v);
#pragma unroll
    for (int i = 0; i < 8; i++) r.v[i] = o[i];
#else
    hostimpl::redc(r.v, t.v);
#endif
}
BN_HD void wide_add(FqWide& a, const FqWide& b) {  // a += b (caller guarantees a+b < 2^512)
#if defined(__CUDA_ARCH__)
    wide_add_ptx(a.v, b.v);
#else
    uint64_t c = 0;
    for (int i = 0; i < 16; i++) {
        uint64_t s = (uint64_t)a.v[i] + b.v[i] + c;
        a.v[i] = (uint32_t)s;
        c = s >> 32;
    }
#endif
}
BN_HD void wide_sub_nb(FqWide& a, const FqWide& b) {  // a -= b for a >= b (no add-back code)
#if defined(__CUDA_ARCH__)
    wide_sub_nb_ptx(a.v, b.v);
#else
    int64_t br = 0;
    for (int i = 0; i < 16; i++) {
        int64_t d = (int64_t)a.v[i] - b.v[i] + br;
        a.v[i] = (uint32_t)d;
        br = d >> 32;
    }
#endif
}
BN_HD void wide_sub(FqWide& a, const FqWide& b) {  // a -= b; on borrow add q*2^256 (stays in [0, qR))
#if defined(__CUDA_ARCH__)
    wide_sub_ptx(a.v, b.v);
#else
    int64_t br = 0;
    for (int i = 0; i < 16; i++) {
        int64_t d = (int64_t)a.v[i] - b.v[i] + br;
        a.v[i] = (uint32_t)d;
        br = d >> 32;
    }
    if (br) {
        uint64_t c = 0;
        for (int i = 0; i < 8; i++) {
            uint64_t s = (uint64_t)a.v[8 + i] + fq_modulus_limb(i) + c;
            a.v[8 + i] = (uint32_t)s;
            c = s >> 32;
        }
    }
#endif
}
// plain 256-bit add without reduction (for Karatsuba sums a0+a1 < 2q < 2^256)
BN_HD void fq_add_noreduce(Fq& r, const Fq& a, const Fq& b) {
#if defined(__CUDA_ARCH__)
    asm("add.cc.u32 %0, %8, %16;\n\taddc.cc.u32 %1, %9, %17;\n\taddc.cc.u32 %2, %10, %18;\n\taddc.cc.u32 %3, %11, %19;\n\t"
        "addc.cc.u32 %4, %12, %20;\n\taddc.cc.u32 %5, %13, %21;\n\taddc.cc.u32 %6, %14, %22;\n\taddc.u32 %7, %15, %23;"
        : "=&r"(r.v[0]), "=&r"(r.v[1]), "=&r"(r.v[2]), "=&r"(r.v[3]), "=&r"(r.v[4]), "=&r"(r.v[5]), "=&r"(r.v[6]), "=&r"(r.v[7])
        : "r"(a.v[0]), "r"(a.v[1]), "r"(a.v[2]), "r"(a.v[3]), "r"(a.v[4]), "r"(a.v[5]), "r"(a.v[6]), "r"(a.v[7]),
          "r"(b.v[0]), "r"(b.v[1]), "r"(b.v[2]), "r"(b.v[3]), "r"(b.v[4]), "r"(b.v[5]), "r"(b.v[6]), "r"(b.v[7]));
#else
    uint64_t c = 0;
    for (int i = 0; i < 8; i++) {
        uint64_t s = (uint64_t)a.v[i] + b.v[i] + c;
        r.v[i] = (uint32_t)s;
        c = s >> 32;
    }
#endif
}

}  // namespace bn
