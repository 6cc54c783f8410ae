// Shared-memory slot file for the hot Fq12 loops (device only).
//
// Measured on B200 (profiles/r01_ncu_l1_l2_traffic.md): the L1 is write-through, so every Fq2 result that an
// out-of-line body stores to local memory travels to the L2 (63 GB per 2^16 final exponentiations, 5 TB/s), and
// with 14 warps per SM sharing ~200 KB of L1 the consumer of that value usually misses (first use of a loaded
// operand: ~400 clk, 60 % of the samples of fq2_add/fq2_sub).  Shared memory is the same SRAM without either
// problem: stores stay on chip and loads always hit.
//
// Each thread owns BN_NSLOTS Fq2 slots (64 B each) in shared memory at a per-thread stride of 464 B = 29 x 16 B:
// an odd number of 16-byte units, so the 8 threads of a quarter-warp hit 8 different bank groups on every 128-bit
// access (conflict-free), and 7 blocks x 64 threads x 464 B = 208 KB fit one SM.  All Fq2 arithmetic of the hot
// loops (cyclotomic squaring and Fq12 multiplication of the final exponentiation's hard part,
// src/fields/fq12.rs:75-105, :198-266, :316-328) runs slot to slot through explicit ld.shared/st.shared bodies;
// the Fq12 values themselves stay in local memory and are touched only by coefficient fetches and result stores.
// Same field elements as the local-memory path in bn_tower.cuh, hence the same bytes.
#pragma once
#include "bn_tower.cuh"

#if defined(__CUDACC__)
namespace bn {

// 7 slots (464-byte stride) for blocks of up to 448 threads, 6 slots (400 bytes = 25 x 16 B, odd as well) for 512-thread blocks:
// 512 x 400 B = 205 KB.  The tower algorithms below use slots 0..5 only; the G2 line steps (bn_slots_curve.cuh) use all 7
// and are launched with at most 448 threads.
#define BN_NSLOTS 7
#define BN_SLOT_STRIDE (BN_NSLOTS * 64 + 16)
#define BN_SLOT_STRIDE6 (6 * 64 + 16)
#define BN_SLOT_BLOCK7 448              /* largest block with 7 slots per thread */
typedef uint32_t sp;   // shared-window byte address of one Fq2 slot

__device__ __forceinline__ uint4 lds128_raw(sp a) {
    uint4 v;
    asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(a) : "memory");
    return v;
}
__device__ __forceinline__ void sts128_raw(sp a, uint4 v) {
    asm volatile("st.shared.v4.u32 [%0], {%1, %2, %3, %4};" :: "r"(a), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}
extern __shared__ uint4 bn_slot_mem[];
// -DBN_SLOT_CHECK (test build `libbn254_b200_chk.so`, scripts/slot_check_cases.py): every shared-memory access of the slot-file
// and warp-cooperative kernels is checked against the region its issuer owns -- the thread's own slots (stride rule of s_base),
// the warp's 32 slot regions in the coalesced result epilogue, the warp's file plus the read-only tables in bn_coop.cuh -- and
// against the block's dynamic shared memory size and the access alignment.  A violation is counted in a device global and the
// access is skipped.  compute-sanitizer is closed on this pool; this is the stand-in for its memcheck, and since every thread
// (warp) then provably touches only memory it owns between barriers, also for racecheck.
#ifdef BN_SLOT_CHECK
__device__ unsigned long long bn_slot_violations[4];   // [0] count, [1] first offending address | kind << 32, [2] blockIdx, [3] threadIdx | blockDim << 32
__device__ __noinline__ void slot_violation(unsigned kind, sp a) {
    if (atomicAdd(&bn_slot_violations[0], 1ull) == 0ull) {
        bn_slot_violations[1] = (unsigned long long)a | ((unsigned long long)kind << 32);
        bn_slot_violations[2] = blockIdx.x;
        bn_slot_violations[3] = threadIdx.x | ((unsigned long long)blockDim.x << 32);
    }
}
// [lo, lo + len) must contain [a, a + bytes), inside the block's dynamic shared memory, naturally aligned
__device__ __forceinline__ bool slot_ok(unsigned kind, sp a, unsigned bytes, sp lo, unsigned len) {
    const sp dyn0 = (sp)__cvta_generic_to_shared(bn_slot_mem);
    unsigned dyn_size;
    asm("mov.u32 %0, %%dynamic_smem_size;" : "=r"(dyn_size));
    const bool ok = (a & (bytes - 1u)) == 0u && a >= lo && a + bytes <= lo + len && lo >= dyn0 && lo + len <= dyn0 + dyn_size;
    if (!ok) slot_violation(kind, a);
    return ok;
}
__device__ __forceinline__ bool slot_ok_own(sp a, unsigned bytes) {
    const unsigned stride = blockDim.x > 448 ? (6 * 64 + 16) : (7 * 64 + 16);
    return slot_ok(1u, a, bytes, (sp)__cvta_generic_to_shared(bn_slot_mem) + threadIdx.x * stride, stride);
}
#define BN_SLOT_GUARD(cond) if (!(cond)) 
#else
#define BN_SLOT_GUARD(cond) if (false)
#endif
// accessors of the slot-file kernels: the calling thread's own slots
__device__ __forceinline__ uint4 lds128(sp a) {
    BN_SLOT_GUARD(slot_ok_own(a, 16u)) return make_uint4(0, 0, 0, 0);
    return lds128_raw(a);
}
__device__ __forceinline__ void sts128(sp a, uint4 v) {
    BN_SLOT_GUARD(slot_ok_own(a, 16u)) return;
    sts128_raw(a, v);
}
__device__ __forceinline__ Fq2 s_ld(sp a) {
    uint4 w0 = lds128(a), w1 = lds128(a + 16), w2 = lds128(a + 32), w3 = lds128(a + 48);
    Fq2 r;
    r.c0.v[0] = w0.x; r.c0.v[1] = w0.y; r.c0.v[2] = w0.z; r.c0.v[3] = w0.w;
    r.c0.v[4] = w1.x; r.c0.v[5] = w1.y; r.c0.v[6] = w1.z; r.c0.v[7] = w1.w;
    r.c1.v[0] = w2.x; r.c1.v[1] = w2.y; r.c1.v[2] = w2.z; r.c1.v[3] = w2.w;
    r.c1.v[4] = w3.x; r.c1.v[5] = w3.y; r.c1.v[6] = w3.z; r.c1.v[7] = w3.w;
    return r;
}
__device__ __forceinline__ void s_st(sp a, const Fq2& v) {
    sts128(a, make_uint4(v.c0.v[0], v.c0.v[1], v.c0.v[2], v.c0.v[3]));
    sts128(a + 16, make_uint4(v.c0.v[4], v.c0.v[5], v.c0.v[6], v.c0.v[7]));
    sts128(a + 32, make_uint4(v.c1.v[0], v.c1.v[1], v.c1.v[2], v.c1.v[3]));
    sts128(a + 48, make_uint4(v.c1.v[4], v.c1.v[5], v.c1.v[6], v.c1.v[7]));
}
__device__ __forceinline__ sp s_base() {
    const unsigned stride = blockDim.x > BN_SLOT_BLOCK7 ? BN_SLOT_STRIDE6 : BN_SLOT_STRIDE;
    return (sp)__cvta_generic_to_shared(bn_slot_mem) + threadIdx.x * stride;
}

// ---- slot-to-slot Fq2 bodies (r may alias a or b) -------------------------------------------------------------
__device__ __noinline__ void s_mul(sp r, sp a, sp b) { Fq2 x = s_ld(a), y = s_ld(b), z; fq2_mul_inl(z, x, y); s_st(r, z); }
__device__ __noinline__ void s_add(sp r, sp a, sp b) { Fq2 x = s_ld(a), y = s_ld(b), z; fq_add(z.c0, x.c0, y.c0); fq_add(z.c1, x.c1, y.c1); s_st(r, z); }
__device__ __noinline__ void s_sub(sp r, sp a, sp b) { Fq2 x = s_ld(a), y = s_ld(b), z; fq_sub(z.c0, x.c0, y.c0); fq_sub(z.c1, x.c1, y.c1); s_st(r, z); }
__device__ __noinline__ void s_dbl(sp r, sp a) { Fq2 x = s_ld(a), z; fq_dbl(z.c0, x.c0); fq_dbl(z.c1, x.c1); s_st(r, z); }
__device__ __noinline__ void s_xi(sp r, sp a) {
    Fq2 x = s_ld(a), n;
    fq_lin9<true>(n.c0, x.c0, x.c1);
    fq_lin9<false>(n.c1, x.c1, x.c0);
    s_st(r, n);
}
// coefficient fetch from local memory: slot = p (mode 0), p + q (1), p - q (2)
__device__ __noinline__ void s_fetch(sp r, const Fq2& p, const Fq2& q, int mode) {
    Fq2 x = fq2_ld(p);
    if (mode != 0) {
        Fq2 y = fq2_ld(q), z;
        if (mode == 1) { fq_add(z.c0, x.c0, y.c0); fq_add(z.c1, x.c1, y.c1); }
        else { fq_sub(z.c0, x.c0, y.c0); fq_sub(z.c1, x.c1, y.c1); }
        x = z;
    }
    s_st(r, x);
}
__device__ __forceinline__ void s_in(sp r, const Fq2& p) { s_st(r, fq2_ld(p)); }
__device__ __forceinline__ void s_out(Fq2& p, sp a) { fq2_st(p, s_ld(a)); }

__device__ __forceinline__ const Fq2& f12c(const Fq12& a, int i) { return reinterpret_cast<const Fq2*>(&a)[i]; }
__device__ __forceinline__ Fq2& f12c(Fq12& a, int i) { return reinterpret_cast<Fq2*>(&a)[i]; }
__device__ __forceinline__ const Fq2& f6c(const Fq6& a, int i) { return reinterpret_cast<const Fq2*>(&a)[i]; }
__device__ __forceinline__ Fq2& f6c(Fq6& a, int i) { return reinterpret_cast<Fq2*>(&a)[i]; }

// ---- Granger-Scott squaring, one coefficient pair per call (value of src/fields/fq12.rs:198-247) --------------
// Fq12 as six Fq2: index 0..5 = c0.c0, c0.c1, c0.c2, c1.c0, c1.c1, c1.c2 = the reference's z0, z4, z3, z2, z1, z5.
//   p = 0: (t0, t1) = sqr4(z0, z1): z0' = 3 t0 - 2 z0, z1' = 3 t1 + 2 z1
//   p = 1: (t2, t3) = sqr4(z2, z3): z4' = 3 t2 - 2 z4, z5' = 3 t3 + 2 z5
//   p = 2: (t4, t5) = sqr4(z4, z5): z3' = 3 t4 - 2 z3, z2' = 3 xi t5 + 2 z2
// r must not overlap a.
// Two register-resident stages around the two products of one pair: the additive work of the Fq4 squaring and of
// the 3t -+ 2z outputs never leaves registers, and the pair costs 4 calls instead of 17.
// pre: slots 0 = a, 1 = b, 3 = S = a + b, 4 = T = a + xi b (operands read from local memory); slot 2 takes TMP = a b
__device__ __noinline__ void s_cyc_pre(sp base, const Fq2& pa, const Fq2& pb) {
    Fq2 a = fq2_ld(pa), b = fq2_ld(pb), s, t;
    s_st(base, a); s_st(base + 64, b);
    fq_add(s.c0, a.c0, b.c0); fq_add(s.c1, a.c1, b.c1);
    s_st(base + 192, s);
    fq_lin9<true>(t.c0, b.c0, b.c1);
    fq_lin9<false>(t.c1, b.c1, b.c0);
    fq_add(t.c0, t.c0, a.c0); fq_add(t.c1, t.c1, a.c1);
    s_st(base + 256, t);
}
// out = 3 t - 2 z (minus) or 3 t + 2 z, every step reduced
__device__ __forceinline__ void fq2_3t2z(Fq2& out, const Fq2& t, const Fq2& z, bool plus) {
    Fq2 d;
    if (plus) { fq_add(d.c0, t.c0, z.c0); fq_add(d.c1, t.c1, z.c1); }
    else { fq_sub(d.c0, t.c0, z.c0); fq_sub(d.c1, t.c1, z.c1); }
    fq_dbl(d.c0, d.c0); fq_dbl(d.c1, d.c1);
    fq_add(out.c0, d.c0, t.c0); fq_add(out.c1, d.c1, t.c1);
}
// post: with S = (a + b)(a + xi b) and TMP = a b in the slots: ta = S - TMP - xi TMP (ta0 = S0 - 10 m0 + m1, ta1 = S1 - 10 m1 - m0),
// tb = 2 TMP (times xi if xi_d); out_c = 3 ta - 2 zc, out_d = 3 tb + 2 zd written to local memory
__device__ __noinline__ void s_cyc_post(Fq2& out_c, Fq2& out_d, const Fq2& zc, const Fq2& zd, sp base, bool xi_d) {
    Fq2 z = fq2_ld(zc), w = fq2_ld(zd);                      // both old coefficients up front: one local-memory latency
    Fq2 m = s_ld(base + 128), s = s_ld(base + 192), u;
    Fq2 oc;
    cyc_outc(oc.c0, s.c0, m.c0, m.c1, true, z.c0);
    cyc_outc(oc.c1, s.c1, m.c1, m.c0, false, z.c1);
    fq2_st(out_c, oc);
    if (xi_d) {
        fq_lin9<true>(u.c0, m.c0, m.c1);
        fq_lin9<false>(u.c1, m.c1, m.c0);
        fq_dbl(m.c0, u.c0); fq_dbl(m.c1, u.c1);              // xi (2 TMP) = 2 (xi TMP)
    } else { fq_dbl(m.c0, m.c0); fq_dbl(m.c1, m.c1); }       // tb
    fq2_3t2z(w, m, w, true);
    fq2_st(out_d, w);
}
__device__ __forceinline__ void s_cyc_pair(Fq12& r, const Fq12& a, int p, sp base) {
    const int ia = (p == 0) ? 0 : (p == 1) ? 3 : 1;
    const int ib = (p == 0) ? 4 : (p == 1) ? 2 : 5;
    const int ic = p;
    const int id = (p == 0) ? 4 : (p == 1) ? 5 : 3;
    s_cyc_pre(base, f12c(a, ia), f12c(a, ib));
    s_mul(base + 128, base, base + 64);             // TMP = a b
    s_mul(base + 192, base + 192, base + 256);      // S = (a + b)(a + xi b)
    s_cyc_post(f12c(r, ic), f12c(r, id), f12c(a, ic), f12c(a, id), base, p == 2);
}
__device__ __forceinline__ void s_cyclotomic_sqr_na(Fq12& r, const Fq12& a, sp base) {
    s_cyc_pair(r, a, 0, base);
    s_cyc_pair(r, a, 1, base);
    s_cyc_pair(r, a, 2, base);
}

// ---- Fq6 product on slots (value of src/fields/fq6.rs:194-208) -------------------------------------------------
#define BN_SKIP 0xffffffffu
// one side of a Karatsuba stage: slots dI <- p_i, dJ <- p_j (skipped when BN_SKIP), dS <- p_i + p_j; operands in local memory
__device__ __noinline__ void s_prep1(const Fq2& pi, const Fq2& pj, sp dI, sp dJ, sp dS) {
    Fq2 x = fq2_ld(pi), y = fq2_ld(pj), s;
    if (dI != BN_SKIP) s_st(dI, x);
    if (dJ != BN_SKIP) s_st(dJ, y);
    fq_add(s.c0, x.c0, y.c0); fq_add(s.c1, x.c1, y.c1);
    s_st(dS, s);
}
// both sides of a stage in one call (all four operands loaded before the first addition: one local-memory latency):
// slots dI <- a_i, dI + 64 <- b_i, dJ <- a_j, dJ + 64 <- b_j (each pair skipped when BN_SKIP), dSA <- a_i + a_j, dSB <- b_i + b_j
__device__ __noinline__ void s_prep2(const Fq2& ai, const Fq2& aj, const Fq2& bi, const Fq2& bj, sp dI, sp dJ, sp dSA, sp dSB) {
    Fq2 x = fq2_ld(ai), y = fq2_ld(aj), u = fq2_ld(bi), w = fq2_ld(bj), s;
    if (dI != BN_SKIP) { s_st(dI, x); s_st(dI + 64, u); }
    if (dJ != BN_SKIP) { s_st(dJ, y); s_st(dJ + 64, w); }
    fq_add(s.c0, x.c0, y.c0); fq_add(s.c1, x.c1, y.c1);
    s_st(dSA, s);
    fq_add(s.c0, u.c0, w.c0); fq_add(s.c1, u.c1, w.c1);
    s_st(dSB, s);
}
// one output coefficient: t = Q - Va - Vb; kind 0: xi t + Vc, kind 1: t + xi Vc, kind 2: t + Vc; then minus (aa + bb) for
// post = 1, minus (aa - bb) for post = 2 (the Fq12-level Karatsuba correction, operands in local memory); result to local memory
__device__ __noinline__ void s_f6_out(Fq2& out, sp Q, sp Va, sp Vb, sp Vc, int kind, const Fq2& aa, const Fq2& bb, int post) {
    if (kind != 2) {
        // the outputs that carry a multiplication by xi: every component is ONE lazily accumulated 288-bit sum, reduced once
        // (|sum| <= 33 q): kind 0  c0 = Vc0 + 9 X0 - X1, c1 = Vc1 + 9 X1 + X0 with X = Q - Va - Vb;  kind 1  c0 = X0 + 9 Vc0 - Vc1,
        // c1 = X1 + 9 Vc1 + Vc0;  then minus (aa +- bb)
        Acc9 x0, x1, t0, t1, v;
        {
            Fq2 a = s_ld(Q), b = s_ld(Va), c = s_ld(Vb);
            acc9_set(x0, a.c0); acc9_subu(x0, b.c0); acc9_subu(x0, c.c0);
            acc9_set(x1, a.c1); acc9_subu(x1, b.c1); acc9_subu(x1, c.c1);
        }
        Fq2 w = s_ld(Vc);
        if (kind == 0) {
            acc9_set(t0, w.c0); acc9_addv(t0, x0); acc9_shl<3>(v, x0); acc9_addv(t0, v); acc9_subv(t0, x1);
            acc9_set(t1, w.c1); acc9_addv(t1, x1); acc9_shl<3>(v, x1); acc9_addv(t1, v); acc9_addv(t1, x0);
        } else {
            Acc9 u;
            t0 = x0; acc9_addu(t0, w.c0); acc9_set(u, w.c0); acc9_shl<3>(v, u); acc9_addv(t0, v); acc9_subu(t0, w.c1);
            t1 = x1; acc9_addu(t1, w.c1); acc9_set(u, w.c1); acc9_shl<3>(v, u); acc9_addv(t1, v); acc9_addu(t1, w.c0);
        }
        if (post) {
            Fq2 x = fq2_ld(aa), y = fq2_ld(bb);
            acc9_subu(t0, x.c0); acc9_subu(t1, x.c1);
            if (post == 1) { acc9_subu(t0, y.c0); acc9_subu(t1, y.c1); }
            else { acc9_addu(t0, y.c0); acc9_addu(t1, y.c1); }
        }
        Fq2 r;
        acc9_reduce<true>(r.c0, t0);
        acc9_reduce<true>(r.c1, t1);
        fq2_st(out, r);
        return;
    }
    // kind 2 (no multiplication by xi): three to five canonical operations are cheaper than a lazy sum and its reduction
    Fq2 t = s_ld(Q), v = s_ld(Va), w = s_ld(Vc), x, y;
    fq_sub(t.c0, t.c0, v.c0); fq_sub(t.c1, t.c1, v.c1);
    v = s_ld(Vb);
    fq_sub(t.c0, t.c0, v.c0); fq_sub(t.c1, t.c1, v.c1);
    x = w; y = t;
    fq_add(t.c0, x.c0, y.c0); fq_add(t.c1, x.c1, y.c1);
    if (post) {
        x = fq2_ld(aa); y = fq2_ld(bb);
        if (post == 1) { fq_add(x.c0, x.c0, y.c0); fq_add(x.c1, x.c1, y.c1); }
        else { fq_sub(x.c0, x.c0, y.c0); fq_sub(x.c1, x.c1, y.c1); }
        fq_sub(t.c0, t.c0, x.c0); fq_sub(t.c1, t.c1, x.c1);
    }
    fq2_st(out, t);
}
// out = a + b (mode 0), a - b (1), a + xi b (2), a - xi b (3), xi b (6); local memory to local memory (the Fq6/Fq12-level
// Karatsuba sums around the slot products)
#define BN_LIN_ADD 0
#define BN_LIN_SUB 1
#define BN_LIN_XI 2
#define BN_LIN_ZERO_A 4
__device__ __noinline__ void l_lin(Fq2& out, const Fq2& a, const Fq2& b, int mode) {
    Fq2 x, y = fq2_ld(b);
    if (mode & BN_LIN_ZERO_A) x = fq2_zero(); else x = fq2_ld(a);
    if (mode & BN_LIN_XI) {
        Fq2 n;
        fq_lin9<true>(n.c0, y.c0, y.c1);
        fq_lin9<false>(n.c1, y.c1, y.c0);
        y = n;
    }
    if (mode & BN_LIN_SUB) { fq_sub(x.c0, x.c0, y.c0); fq_sub(x.c1, x.c1, y.c1); }
    else { fq_add(x.c0, x.c0, y.c0); fq_add(x.c1, x.c1, y.c1); }
    fq2_st(out, x);
}
// The same over the three coefficients of an Fq6 in one call: out = a +- b, or a +- v b with rot (v b = (xi b2, b0, b1),
// src/fields/fq6.rs:109-115).  The operands of the next coefficient are loaded before the current one is computed, so the
// call exposes one local-memory latency instead of three.  out must not overlap a or b.
__device__ __noinline__ void l_lin6(Fq6& out, const Fq6& a, const Fq6& b, int mode, bool rot) {
    Fq2 x = fq2_ld(a.c0), y = fq2_ld(rot ? b.c2 : b.c0);
#pragma unroll 1
    for (int i = 0; i < 3; i++) {
        const int n = i < 2 ? i + 1 : 2;
        Fq2 xn = fq2_ld(f6c(a, n)), yn = fq2_ld(f6c(b, rot ? n - 1 : n));
        if (rot && i == 0) {
            Fq2 t;
            fq_lin9<true>(t.c0, y.c0, y.c1);
            fq_lin9<false>(t.c1, y.c1, y.c0);
            y = t;
        }
        if (mode & BN_LIN_SUB) { fq_sub(x.c0, x.c0, y.c0); fq_sub(x.c1, x.c1, y.c1); }
        else { fq_add(x.c0, x.c0, y.c0); fq_add(x.c1, x.c1, y.c1); }
        fq2_st(f6c(out, i), x);
        x = xn; y = yn;
    }
}
// out = a * b, operands and result in local memory, out not overlapping them.  post != 0 finishes an Fq12-level Karatsuba
// product in the same pass: out_i = (a b)_i - (p_i + r_i) for post = 1, (a b)_i - (p_i - r_i) for post = 2.
// Karatsuba (6 products) in stages of 3 + 1 + 1 + 1 products on 6 slots (the first output leaves before the last product); 15 calls.
__device__ __noinline__ void s_fq6_mul_p(Fq6& out, const Fq6& a, const Fq6& b, const Fq2& p0, const Fq2& r0, const Fq2& p1, const Fq2& r1,
                                         const Fq2& p2, const Fq2& r2, int post, sp base) {
    const sp s0 = base, s1 = base + 64, s2 = base + 128, s3 = base + 192, s4 = base + 256, s5 = base + 320;
    s_prep2(a.c0, a.c1, b.c0, b.c1, s0, s2, s4, s5);
    s_mul(s0, s0, s1);                       // v0 = a0 b0
    s_mul(s2, s2, s3);                       // v1 = a1 b1
    s_mul(s4, s4, s5);                       // (a0 + a1)(b0 + b1)
    s_in(s1, a.c2); s_in(s3, b.c2);
    s_mul(s1, s1, s3);                       // v2 = a2 b2
    s_prep2(a.c1, a.c2, b.c1, b.c2, BN_SKIP, BN_SKIP, s3, s5);
    s_mul(s3, s3, s5);                       // (a1 + a2)(b1 + b2)
    s_f6_out(out.c0, s3, s2, s1, s0, 0, p0, r0, post);    // v0 + xi ((a1+a2)(b1+b2) - v1 - v2); frees slot 3
    s_prep2(a.c0, a.c2, b.c0, b.c2, BN_SKIP, BN_SKIP, s3, s5);
    s_mul(s3, s3, s5);                       // (a0 + a2)(b0 + b2)
    s_f6_out(out.c1, s4, s0, s2, s1, 1, p1, r1, post);    // (a0+a1)(b0+b1) - v0 - v1 + xi v2
    s_f6_out(out.c2, s3, s0, s1, s2, 2, p2, r2, post);    // (a0+a2)(b0+b2) - v0 - v2 + v1
}
__device__ __forceinline__ void s_fq6_mul(Fq6& out, const Fq6& a, const Fq6& b, const Fq6& aa, const Fq6& bb, int post, sp base) {
    s_fq6_mul_p(out, a, b, aa.c0, bb.c0, aa.c1, bb.c1, aa.c2, bb.c2, post, base);
}
// r = a * b (conj_b: a * conj(b)); value of src/fields/fq12.rs:316-328.  r must not overlap a or b.
__device__ __noinline__ void s_fq12_mul_na(Fq12& r, const Fq12& a, const Fq12& b, bool conj_b, sp base) {
    Fq6 aa, bb, sb;
    const int m = conj_b ? BN_LIN_SUB : BN_LIN_ADD;
    s_fq6_mul(aa, a.c0, b.c0, aa, aa, 0, base);
    s_fq6_mul(bb, a.c1, b.c1, bb, bb, 0, base);
    // r.c0 doubles as the buffer of a0 + a1; c1 = (a0 + a1)(b0 +- b1) - (aa +- bb)
    l_lin6(r.c0, a.c0, a.c1, BN_LIN_ADD, false);
    l_lin6(sb, b.c0, b.c1, m, false);
    s_fq6_mul(r.c1, r.c0, sb, aa, bb, conj_b ? 2 : 1, base);
    // c0 = aa +- v bb, v (b0, b1, b2) = (xi b2, b0, b1)
    l_lin6(r.c0, aa, bb, m, true);
}
// r = a^2, complex method (value of src/fields/fq12.rs:295-303): ab = a0 a1, c0 = (a0 + a1)(a0 + v a1) - ab - v ab, c1 = 2 ab.
// r must not overlap a.
__device__ __noinline__ void s_fq12_sqr_na(Fq12& r, const Fq12& a, sp base) {
    Fq6 ab, sv;
    Fq2 xab2;
    s_fq6_mul(ab, a.c0, a.c1, ab, ab, 0, base);
    // r.c1 doubles as the buffer of a0 + a1
    l_lin6(r.c1, a.c0, a.c1, BN_LIN_ADD, false);
    l_lin6(sv, a.c0, a.c1, BN_LIN_ADD, true);
    l_lin(xab2, ab.c2, ab.c2, BN_LIN_XI | BN_LIN_ZERO_A);
    s_fq6_mul_p(r.c0, r.c1, sv, ab.c0, xab2, ab.c1, ab.c0, ab.c2, ab.c1, 1, base);
    l_lin6(r.c1, ab, ab, BN_LIN_ADD, false);
}

// ---- line multiplication as six 3-term dot products (value of src/fields/fq12.rs:130-196) ----------------------
// out = sum_{k<3} z[zi_k] * x[xi_k] with one Montgomery reduction per component: the three Fq2 products are accumulated
// unreduced (each Karatsuba: z0 x0, z1 x1, (z0 + z1)(x0 + x1); sums < 12 q^2 < 2^512), then
// c0 = sum z0 x0 - sum z1 x1 (+ q 2^256 when negative), c1 = sum (z0+z1)(x0+x1) - sum z0 x0 - sum z1 x1 < 6 q^2 (one
// conditional subtraction of q 2^256 brings it below q 2^256).  z from the Fq12 in local memory (3-bit indices packed in
// zi), x from the slots xbase + 64 * index (packed in xi).  Same IMAD work as the 13-product Karatsuba form, no additive glue.
__device__ __noinline__ void s_dot3(Fq2& out, const Fq12& f, uint32_t zi, sp xbase, uint32_t xi) {
    FqWide t0, t1, s;
    Fq2 z = fq2_ld(f12c(f, zi & 7));
    {   // first term: plain products
        Fq2 x = s_ld(xbase + 64 * (xi & 7));
        Fq2 zn = fq2_ld(f12c(f, (zi >> 3) & 7));      // next coefficient: its local-memory latency rides under the products
        fq_mulwide(t0, z.c0, x.c0);
        fq_mulwide(t1, z.c1, x.c1);
        Fq sa, sb;
        fq_add_noreduce(sa, z.c0, z.c1);
        fq_add_noreduce(sb, x.c0, x.c1);
        fq_mulwide(s, sa, sb);
        z = zn;
    }
#pragma unroll 1
    for (int k = 1; k < 3; k++) {
        Fq2 x = s_ld(xbase + 64 * ((xi >> (3 * k)) & 7));
        Fq2 zn = fq2_ld(f12c(f, (zi >> (3 * (k < 2 ? k + 1 : k))) & 7));
        FqWide p;
        fq_mulwide(p, z.c0, x.c0); wide_add(t0, p);
        fq_mulwide(p, z.c1, x.c1); wide_add(t1, p);
        Fq sa, sb;
        fq_add_noreduce(sa, z.c0, z.c1);
        fq_add_noreduce(sb, x.c0, x.c1);
        fq_mulwide(p, sa, sb); wide_add(s, p);
        z = zn;
    }
    wide_sub_nb(s, t0);                 // never borrows
    wide_sub_nb(s, t1);
    wide_sub(t0, t1);                // + q 2^256 on borrow: in [0, q 2^256)
    Fq hi, zero = fq_zero();
#pragma unroll
    for (int i = 0; i < 8; i++) hi.v[i] = s.v[8 + i];
    fq_add(hi, hi, zero);            // high half < 1.14 q: one conditional subtraction of q
#pragma unroll
    for (int i = 0; i < 8; i++) s.v[8 + i] = hi.v[i];
    Fq2 r;
    fq_redc(r.c0, t0);
    fq_redc(r.c1, s);
    fq2_st(out, r);
}
// slots d <- c * k, dxi <- xi c k for the Fq2 coefficient c read from the line table (four 16-byte words, stride words apart):
// ell_vv.scale(P.x) / ell_vw.scale(P.y) of src/groups/mod.rs:589
__device__ __noinline__ void s_scale_xi(sp d, sp dxi, const uint4* __restrict__ g, size_t stride, const Fq& k) {
    uint4 w0 = __ldg(g), w1 = __ldg(g + stride), w2 = __ldg(g + 2 * stride), w3 = __ldg(g + 3 * stride);
    Fq kk = fq_ld(k);
    Fq2 c, r, n;
    c.c0.v[0] = w0.x; c.c0.v[1] = w0.y; c.c0.v[2] = w0.z; c.c0.v[3] = w0.w; c.c0.v[4] = w1.x; c.c0.v[5] = w1.y; c.c0.v[6] = w1.z; c.c0.v[7] = w1.w;
    c.c1.v[0] = w2.x; c.c1.v[1] = w2.y; c.c1.v[2] = w2.z; c.c1.v[3] = w2.w; c.c1.v[4] = w3.x; c.c1.v[5] = w3.y; c.c1.v[6] = w3.z; c.c1.v[7] = w3.w;
    fq_mul(r.c0, c.c0, kk);
    fq_mul(r.c1, c.c1, kk);
    s_st(d, r);
    fq_lin9<true>(n.c0, r.c0, r.c1);
    fq_lin9<false>(n.c1, r.c1, r.c0);
    s_st(dxi, n);
}
// r = f * (ell_0 + ell_vv P.x v^2... ) i.e. one line evaluated at P and multiplied into f (line_mul of bn_curve.cuh).
// With f = z0 + z3 w + z1 w^2 + z4 w^3 + z2 w^4 + z5 w^5 (w^6 = xi) and the line x0 + x4 w^3 + x2 w^4:
//   z0' = z0 x0 + xi (z1 x2 + z4 x4)      z3' = z3 x0 + xi (z2 x4 + z4 x2)
//   z1' = z1 x0 + xi (z5 x4 + z2 x2)      z4' = z4 x0 + z0 x4 + xi z5 x2
//   z2' = z2 x0 + z0 x2 + z3 x4           z5' = z5 x0 + z3 x2 + z1 x4
// Slots: 0 = x0, 1 = x2, 2 = x4, 3 = xi x2, 4 = xi x4.  r must not overlap f.
#define BN_PK3(a, b, c) ((uint32_t)(a) | ((uint32_t)(b) << 3) | ((uint32_t)(c) << 6))
__device__ __noinline__ void s_line_mul(Fq12& r, const Fq12& f, const uint4* __restrict__ c, size_t nt, const AffG1& p, sp base) {
    sts128(base, __ldg(c)); sts128(base + 16, __ldg(c + nt)); sts128(base + 32, __ldg(c + 2 * nt)); sts128(base + 48, __ldg(c + 3 * nt));
    s_scale_xi(base + 64, base + 192, c + 8 * nt, nt, p.x);      // x2 = ell_vv P.x
    s_scale_xi(base + 128, base + 256, c + 4 * nt, nt, p.y);     // x4 = ell_vw P.y
    s_dot3(r.c0.c0, f, BN_PK3(0, 1, 4), base, BN_PK3(0, 3, 4));
    s_dot3(r.c0.c1, f, BN_PK3(1, 5, 2), base, BN_PK3(0, 4, 3));
    s_dot3(r.c0.c2, f, BN_PK3(2, 0, 3), base, BN_PK3(0, 1, 2));
    s_dot3(r.c1.c0, f, BN_PK3(3, 2, 4), base, BN_PK3(0, 4, 3));
    s_dot3(r.c1.c1, f, BN_PK3(4, 0, 5), base, BN_PK3(0, 2, 3));
    s_dot3(r.c1.c2, f, BN_PK3(5, 3, 1), base, BN_PK3(0, 1, 2));
}

// conj(a^u): same width-4 signed-window chain as fq12_exp_by_neg_z (bn_tower.cuh), on the slot bodies
// SYNC: the whole block (one per SM) changes between the squaring code and the multiplication code together, so the
// hot SASS of either phase -- not their union -- has to fit the 32 KB instruction cache (every thread runs the same
// fixed schedule: the exponent is a constant).
template <bool SYNC> __device__ __forceinline__ void s_phase() { if (SYNC) __syncthreads(); }
template <bool SYNC> __device__ __noinline__ void s_fq12_exp_by_neg_z(Fq12& r, const Fq12& a, sp base) {
    Fq12 a3, a5, a7, buf0, buf1;
    s_cyclotomic_sqr_na(buf0, a, base);
    s_phase<SYNC>();
    s_fq12_mul_na(a3, buf0, a, false, base);
    s_fq12_mul_na(a5, buf0, a3, false, base);
    s_fq12_mul_na(a7, buf0, a5, false, base);
    s_phase<SYNC>();
    bool in0 = true;
    for (int i = 61; i >= 0; i--) {
        if (i != 61) {
            if (in0) s_cyclotomic_sqr_na(buf1, buf0, base); else s_cyclotomic_sqr_na(buf0, buf1, base);
            in0 = !in0;
        }
        if ((BN_U_WNAF4_NZ >> i) & 1) {
            bool neg = ((BN_U_WNAF4_NEG >> i) & 1) != 0;
            int idx = (int)((BN_U_WNAF4_I0 >> i) & 1) | ((int)((BN_U_WNAF4_I1 >> i) & 1) << 1);
            s_phase<SYNC>();
            if (idx == 0) { if (in0) s_fq12_mul_na(buf1, buf0, a, neg, base); else s_fq12_mul_na(buf0, buf1, a, neg, base); }
            else if (idx == 1) { if (in0) s_fq12_mul_na(buf1, buf0, a3, neg, base); else s_fq12_mul_na(buf0, buf1, a3, neg, base); }
            else if (idx == 2) { if (in0) s_fq12_mul_na(buf1, buf0, a5, neg, base); else s_fq12_mul_na(buf0, buf1, a5, neg, base); }
            else { if (in0) s_fq12_mul_na(buf1, buf0, a7, neg, base); else s_fq12_mul_na(buf0, buf1, a7, neg, base); }
            s_phase<SYNC>();
            in0 = !in0;
        }
    }
    if (in0) fq12_conj(r, buf0); else fq12_conj(r, buf1);
}
// hard part, the reference's addition chain (src/fields/fq12.rs:75-105, step names in the comments); every product
// writes a buffer that is dead at that point, so the non-aliasing slot bodies apply.  out must not overlap x.
template <bool SYNC> __device__ __noinline__ void s_fq12_final_exp_last_chunk(Fq12& out, const Fq12& x, sp base) {
    Fq12 a, b, c, d, e, f, g;
    s_fq12_exp_by_neg_z<SYNC>(a, x, base);    // a
    s_cyclotomic_sqr_na(b, a, base);          // b
    s_cyclotomic_sqr_na(c, b, base);          // c
    s_phase<SYNC>();
    s_fq12_mul_na(d, c, b, false, base);      // d
    s_phase<SYNC>();
    s_fq12_exp_by_neg_z<SYNC>(e, d, base);    // e
    s_cyclotomic_sqr_na(f, e, base);          // f
    s_fq12_exp_by_neg_z<SYNC>(g, f, base);    // g
    s_phase<SYNC>();
    s_fq12_mul_na(a, e, g, true, base);       // j = conj(g) e
    s_fq12_mul_na(c, a, d, true, base);       // k = j conj(d)
    s_fq12_mul_na(f, c, b, false, base);      // l = k b
    s_fq12_mul_na(a, c, e, false, base);      // m = k e
    s_fq12_mul_na(g, x, a, false, base);      // n = x m
    s_phase<SYNC>();
    fq12_frobenius(a, f, 1);                  // o
    s_phase<SYNC>();
    s_fq12_mul_na(b, a, g, false, base);      // p = o n
    s_phase<SYNC>();
    fq12_frobenius(a, c, 2);                  // q
    s_phase<SYNC>();
    s_fq12_mul_na(g, a, b, false, base);      // r = q p
    s_fq12_mul_na(a, f, x, true, base);       // t = conj(x) l
    s_phase<SYNC>();
    fq12_frobenius(b, a, 3);                  // u
    s_phase<SYNC>();
    s_fq12_mul_na(out, b, g, false, base);    // v = u r
}

}  // namespace bn
#endif
