// Quad-lane executor of the cooperative pairing microcode: the lowest-latency path (BASELINE configs[0], src/lib.rs:611-613 at
// n = 1; examples/joux.rs:17-21).
//
// bn_coop.cuh gives every lane of ONE warp a whole Fq2 product per MUL step: a 336-product carry-chained sequence on one
// sub-partition's multiplier, 2.74 ms for a pairing (profiles/r02_ncu_full_coop.md: one warp issues an instruction every
// 3.7 clk).  Here the SAME tables (bn_coop_ucode.cuh, proven on the CPU by tools/gen_coop_ucode.py) run on a GROUP of four
// warps, four threads per lane of the microcode:
//     MUL  thread 0: x0 y0, 1: x1 y1, 2: x0 y1, 3: x1 y0 (one 8x8-limb product each, operands accumulated per Fq component);
//          the odd threads hand their product to the even ones by shuffle; thread 0 reduces x0 y0 - x1 y1, thread 2 reduces
//          x0 y1 + x1 y0: 64 + 72 multiplier instructions in sequence instead of 336
//     LIN  thread 0 produces c0 = P0 + 9 X0 - X1, thread 2 produces c1 = P1 + 9 X1 + X0
//     INV  thread 0 of the lane
// Lane v of the microcode lives on warp v / 8, so the four warps' multipliers (one per SM sub-partition) work side by side; a
// step ends with a named barrier over the group's 128 threads.  Same field elements as every other path, hence the same bytes.
#pragma once
#include "bn_coop.cuh"

#if defined(__CUDACC__)
namespace bn {

#define BN_COOP4_GROUP 128u
#define BN_COOP4_MAX_GROUPS 4                                  /* groups (pairings) per block: 57 KB of tables + 4 x 11,136 B */
#define BN_COOP4_CODE_WORDS (BN_COOP_NSTEPS * 160)                /* every step: the sequential routines and the paired ones */
#define BN_COOP4_TABLE_WORDS ((BN_COOP4_CODE_WORDS + BN_COOP_NSTEPS + 2 * BN_COOP4_NTRACE + 3) & ~3)
#define BN_COOP4_SMEM(groups) ((size_t)BN_COOP4_TABLE_WORDS * 4 + (size_t)(groups) * BN_COOP_SLOT_BYTES)
#define BN_COOP4_DEFAULT_MAX_N 592                             /* 148 SMs x 4 groups */

#ifdef BN_SLOT_CHECK
__device__ __forceinline__ bool coop4_ok(sp a, unsigned bytes, bool write) {
    const sp dyn0 = (sp)__cvta_generic_to_shared(bn_slot_mem);
    const sp file = dyn0 + 4u * BN_COOP4_TABLE_WORDS + (threadIdx.x / BN_COOP4_GROUP) * BN_COOP_SLOT_BYTES;
    if (!write && a < dyn0 + 4u * BN_COOP4_TABLE_WORDS) return slot_ok(6u, a, bytes, dyn0, 4u * BN_COOP4_TABLE_WORDS);
    return slot_ok(write ? 8u : 7u, a, bytes, file, BN_COOP_SLOT_BYTES);
}
#endif
__device__ __forceinline__ uint32_t q_lds32(sp a) {
    uint32_t v = 0;
    BN_SLOT_GUARD(coop4_ok(a, 4u, false)) return v;
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(a) : "memory");
    return v;
}
__device__ __forceinline__ void q_sts32(sp a, uint32_t v) {
    BN_SLOT_GUARD(coop4_ok(a, 4u, true)) return;
    asm volatile("st.shared.u32 [%0], %1;" :: "r"(a), "r"(v) : "memory");
}
__device__ __forceinline__ uint4 q_lds128(sp a) {
    BN_SLOT_GUARD(coop4_ok(a, 16u, false)) return make_uint4(0, 0, 0, 0);
    return lds128_raw(a);
}
__device__ __forceinline__ void q_sts128(sp a, uint4 v) {
    BN_SLOT_GUARD(coop4_ok(a, 16u, true)) return;
    sts128_raw(a, v);
}
// one Fq component (32 bytes) of a slot
__device__ __forceinline__ Fq q_ld_fq(sp a) {
    uint4 w0 = q_lds128(a), w1 = q_lds128(a + 16);
    Fq r;
    r.v[0] = w0.x; r.v[1] = w0.y; r.v[2] = w0.z; r.v[3] = w0.w; r.v[4] = w1.x; r.v[5] = w1.y; r.v[6] = w1.z; r.v[7] = w1.w;
    return r;
}
__device__ __forceinline__ void q_st_fq(sp a, const Fq& v) {
    q_sts128(a, make_uint4(v.v[0], v.v[1], v.v[2], v.v[3]));
    q_sts128(a + 16, make_uint4(v.v[4], v.v[5], v.v[6], v.v[7]));
}
__device__ __forceinline__ sp coop4_slot(uint32_t t, sp wbase, uint32_t bases) {
    return wbase + ((t & 0xffu) + ((bases >> (8u * ((t >> 8) & 3u))) & 0xffu)) * 64u;
}
// term j (0..7) of a lane's queue; the sign of a term as an all-ones mask
__device__ __forceinline__ uint32_t coop4_term(unsigned long long tq_lo, unsigned long long tq_hi, unsigned j) {
    return (uint32_t)((j & 4u ? tq_hi : tq_lo) >> (16u * (j & 3u))) & 0xffffu;
}
__device__ __forceinline__ uint32_t coop4_sign(uint32_t t) { return 0u - ((t >> 10) & 1u); }
__device__ __forceinline__ void coop4_sync(unsigned group) { asm volatile("bar.sync %0, 128;" :: "r"(group + 1u) : "memory"); }

// One step of the microcode for thread `part` (0..3) of lane `vlane` (decode as coop_step, bn_coop.cuh).  Operand sums are
// accumulated WITHOUT a reduction per term (Acc9, bn_fq.cuh) and reduced once; the two sums of a step advance side by side in
// one loop (independent carry chains); a missing term is slot 0 (the constant zero).
__device__ __forceinline__ void coop4_step(uint32_t tr, sp s_code, sp s_hdr, sp wbase, unsigned vlane, unsigned part) {
    const uint32_t step = tr & 0xffu, bases = tr & 0xffffff00u;
    const uint32_t hdr = q_lds32(s_hdr + 4u * step);
    const sp cw = s_code + 4u * (step * 160u + vlane);
    const uint32_t w0 = q_lds32(cw);
    const unsigned long long tq_lo = (unsigned long long)q_lds32(cw + 128u) | ((unsigned long long)q_lds32(cw + 256u) << 32);
    const unsigned long long tq_hi = (unsigned long long)q_lds32(cw + 384u) | ((unsigned long long)q_lds32(cw + 512u) << 32);
    const unsigned op = hdr & 15u, n0 = (hdr >> 4) & 15u, n1 = (hdr >> 8) & 15u;
    const unsigned nmax = n0 > n1 ? n0 : n1;
    const bool active = (w0 >> 10) & 1u;
    if (!__any_sync(0xffffffffu, active)) return;                // a warp none of whose eight lanes takes part waits at the barrier
    const sp dst = coop4_slot(w0, wbase, bases);
    if (op == 1u) {                                              // MUL: every thread one Fq x Fq product
        const unsigned cx = part & 1u, cy = (part ^ (part >> 1)) & 1u;      // part 0: x0 y0, 1: x1 y1, 2: x0 y1, 3: x1 y0
        // conjugation negates the c1 component: folded into the signs of the terms
        const uint32_t fx = ((w0 & 0x800u) && cx) ? 0xffffffffu : 0u, fy = ((w0 & 0x1000u) && cy) ? 0xffffffffu : 0u;
        Acc9 ax, ay;
        acc9_zero(ax); acc9_zero(ay);
        // the operands of term k + 1 are fetched before term k is added (one shared-memory latency per step, not per term)
        uint32_t ta = n0 ? coop4_term(tq_lo, tq_hi, 0u) : 0u, tb = n1 ? coop4_term(tq_lo, tq_hi, n0) : 0u;
        Fq va = q_ld_fq(coop4_slot(ta, wbase, bases) + 32u * cx), vb = q_ld_fq(coop4_slot(tb, wbase, bases) + 32u * cy);
#pragma unroll 1
        for (unsigned k = 0; k < nmax; k++) {
            const uint32_t na = k + 1u < n0 ? coop4_term(tq_lo, tq_hi, k + 1u) : 0u, nb = k + 1u < n1 ? coop4_term(tq_lo, tq_hi, n0 + k + 1u) : 0u;
            const Fq wa = q_ld_fq(coop4_slot(na, wbase, bases) + 32u * cx), wb = q_ld_fq(coop4_slot(nb, wbase, bases) + 32u * cy);
            acc9_add(ax, va, coop4_sign(ta) ^ fx);
            acc9_add(ay, vb, coop4_sign(tb) ^ fy);
            ta = na; tb = nb; va = wa; vb = wb;
        }
        Fq x, y;
        acc9_reduce<false>(x, ax);
        acc9_reduce<false>(y, ay);
        FqWide t, u;
        fq_mulwide(t, x, y);
#pragma unroll
        for (int k = 0; k < 16; k++) u.v[k] = __shfl_down_sync(0xffffffffu, t.v[k], 1);
        if (!(part & 1u)) {
            if (part == 0u) wide_sub(t, u);                      // x0 y0 - x1 y1 (+ q 2^256 when negative)
            else wide_add(t, u);                                 // x0 y1 + x1 y0 < 2 q^2
            Fq r;
            fq_redc(r, t);
            if (active) q_st_fq(dst + 16u * part, r);            // part 0 -> c0 (offset 0), part 2 -> c1 (offset 32)
        }
    } else if (op == 2u) {                                       // LIN: threads 0 and 2 one component each
        const unsigned comp = part >> 1;
        Fq r = fq_zero();
        if (!(part & 1u)) {
            // c0 = P0 + 9 X0 - X1, c1 = P1 + 9 X1 + X0: the cross component joins the plain sum, the own one is scaled by 9
            const uint32_t fc = comp ? 0u : 0xffffffffu;
            Acc9 ap, a9;
            acc9_zero(ap); acc9_zero(a9);
            uint32_t tp = n0 ? coop4_term(tq_lo, tq_hi, 0u) : 0u, tx = n1 ? coop4_term(tq_lo, tq_hi, n0) : 0u;
            sp sx = coop4_slot(tx, wbase, bases);
            Fq vp = q_ld_fq(coop4_slot(tp, wbase, bases) + 32u * comp), vo = q_ld_fq(sx + 32u * comp), vc = q_ld_fq(sx + 32u * (comp ^ 1u));
#pragma unroll 1
            for (unsigned k = 0; k < nmax; k++) {
                const uint32_t np = k + 1u < n0 ? coop4_term(tq_lo, tq_hi, k + 1u) : 0u, nx = k + 1u < n1 ? coop4_term(tq_lo, tq_hi, n0 + k + 1u) : 0u;
                sx = coop4_slot(nx, wbase, bases);
                const Fq wp = q_ld_fq(coop4_slot(np, wbase, bases) + 32u * comp), wo = q_ld_fq(sx + 32u * comp), wc = q_ld_fq(sx + 32u * (comp ^ 1u));
                acc9_add(ap, vp, coop4_sign(tp));
                acc9_add(a9, vo, coop4_sign(tx));
                acc9_add(ap, vc, coop4_sign(tx) ^ fc);
                tp = np; tx = nx; vp = wp; vo = wo; vc = wc;
            }
            if (n1) { acc9_add9(ap, a9); acc9_reduce<true>(r, ap); }
            else acc9_reduce<false>(r, ap);
            if (hdr & 0x2000u) {
                Fq h;
                fq_half_r(h, r);
                if (w0 & 0x2000u) r = h;
            }
        }
        __syncwarp();                                            // in-place lanes: both components read before either is written
        if (active && !(part & 1u)) q_st_fq(dst + 32u * comp, r);
    } else {                                                     // INV: thread 0 of the one active lane
        if (active && part == 0u) {
            Acc9 ax;
            acc9_zero(ax);
#pragma unroll 1
            for (unsigned k = 0; k < n0; k++) {
                const uint32_t t = coop4_term(tq_lo, tq_hi, k);
                acc9_add(ax, q_ld_fq(coop4_slot(t, wbase, bases)), coop4_sign(t));
            }
            Fq x, r;
            acc9_reduce<false>(x, ax);
            fq_inv_bgcd(r, x);
            q_st_fq(dst, r);
            q_st_fq(dst + 32u, fq_zero());
        }
    }
}
__device__ __forceinline__ void coop4_call(unsigned begin, unsigned count, uint32_t bases, sp s_code, sp s_hdr, sp wbase, unsigned vlane,
                                           unsigned part, unsigned group) {
#pragma unroll 1
    for (unsigned k = 0; k < count; k++) {
        coop4_step((begin + k) | bases, s_code, s_hdr, wbase, vlane, part);
        coop4_sync(group);
    }
}

// the whole step table, its headers and the paired trace travel to shared memory once per block (57 KB)
__device__ __forceinline__ sp coop4_load_tables() {
    uint32_t* dst = reinterpret_cast<uint32_t*>(bn_slot_mem);
    for (unsigned k = threadIdx.x; k < BN_COOP4_CODE_WORDS; k += blockDim.x) dst[k] = bn_coop_code[k];
    for (unsigned k = threadIdx.x; k < BN_COOP_NSTEPS; k += blockDim.x) dst[BN_COOP4_CODE_WORDS + k] = bn_coop_hdr[k];
    for (unsigned k = threadIdx.x; k < 2u * BN_COOP4_NTRACE; k += blockDim.x) dst[BN_COOP4_CODE_WORDS + BN_COOP_NSTEPS + k] = bn_coop4_trace[k];
    __syncthreads();
    return (sp)__cvta_generic_to_shared(bn_slot_mem);
}

// Modes, arguments and results of k_coop_pairing (bn_coop.cuh); `groups` pairings per block, 128 threads each.
__global__ void __launch_bounds__(BN_COOP4_MAX_GROUPS * BN_COOP4_GROUP, 1) k_coop4_pairing(const void* g1_or_gt, const G2* __restrict__ g2,
                                  Fq12* out, uint8_t* __restrict__ ok, uint8_t* __restrict__ is_one, size_t n, int mode) {
    const sp tab = coop4_load_tables();
    const sp s_code = tab, s_hdr = tab + 4u * BN_COOP4_CODE_WORDS, s_trace = s_hdr + 4u * BN_COOP_NSTEPS;
    const unsigned group = threadIdx.x / BN_COOP4_GROUP, gt = threadIdx.x % BN_COOP4_GROUP;
    const unsigned vlane = ((((gt >> 5) + group) & 3u) << 3) | ((gt >> 2) & 7u), part = gt & 3u;   // lanes 0..7 (the busiest) of the block's groups on different sub-partitions
    const size_t i = (size_t)blockIdx.x * (blockDim.x / BN_COOP4_GROUP) + group;
    if (i >= n) return;                                          // whole groups leave: the barriers below are per group
    const sp wbase = tab + 4u * BN_COOP4_TABLE_WORDS + group * BN_COOP_SLOT_BYTES;

    for (unsigned k = gt; k < BN_COOP_NCONST * 16u; k += BN_COOP4_GROUP) q_sts32(wbase + 4u * k, bn_coop_const[k]);
    bool good = true;
    int t_lo, t_hi;
    if (mode == 2) {
        const uint32_t* src = reinterpret_cast<const uint32_t*>(reinterpret_cast<const Fq12*>(g1_or_gt) + i);
        uint32_t any = 0;
        for (unsigned k = 0; k < 96u; k++) any |= src[k];        // every thread the same words: a uniform decision
        for (unsigned k = gt; k < 96u; k += BN_COOP4_GROUP) q_sts32(wbase + 64u * BN_COOP_SLOT_F + 4u * k, src[k]);
        good = any != 0;                                         // zero has no final exponentiation (Option::None)
        t_lo = BN_COOP4_T_FEXP_BEGIN; t_hi = BN_COOP4_T_FEXP_END;
    } else {
        const uint32_t* p = reinterpret_cast<const uint32_t*>(reinterpret_cast<const G1*>(g1_or_gt) + i);
        const uint32_t* q = reinterpret_cast<const uint32_t*>(g2 + i);
        for (unsigned k = gt; k < 48u; k += BN_COOP4_GROUP) {
            const unsigned c = k >> 4, w = k & 15u;
            q_sts32(wbase + 64u * (BN_COOP_SLOT_IN_P + c) + 4u * w, w < 8u ? p[8u * c + w] : 0u);
        }
        for (unsigned k = gt; k < 48u; k += BN_COOP4_GROUP) q_sts32(wbase + 64u * BN_COOP_SLOT_IN_Q + 4u * k, q[k]);
        uint32_t pz = 0, qz = 0;
        for (unsigned w = 0; w < 8u; w++) pz |= p[16u + w];
        for (unsigned w = 0; w < 16u; w++) qz |= q[32u + w];
        good = pz != 0 && qz != 0;
        t_lo = BN_COOP4_T_PAIRING_BEGIN;
        t_hi = mode == 1 ? BN_COOP4_T_MILLER_END : BN_COOP4_T_PAIRING_END;
    }
    coop4_sync(group);

    if (good) {
#pragma unroll 1
        for (int t = t_lo; t < t_hi; t++) {
            // a paired entry: lanes 24..31 (one warp of the group) run lanes 0..7 of step W, the G2 line walk, beside step F
            const uint32_t w0 = q_lds32(s_trace + 8u * (unsigned)t), w1 = q_lds32(s_trace + 8u * (unsigned)t + 4u);
            uint32_t sid = w0 & 0xffu, ln = vlane;
            const uint32_t sw = (w0 >> 8) & 0xffu;
            if (sw != 0xffu && vlane >= BN_COOP4_W_LANE0) { sid = sw; ln = vlane - BN_COOP4_W_LANE0; }
            if (sid != 0xffu) coop4_step(sid | (w1 & 0xffffff00u), s_code, s_hdr, wbase, ln, part);
            coop4_sync(group);
        }
    }
    if (gt >= 32u) return;                                       // the group's first warp writes the result
    const unsigned lane = gt;
    const Fq one = fq_one();
    uint4 one_w = make_uint4(0, 0, 0, 0);
    if (lane == 0) one_w = make_uint4(one.v[0], one.v[1], one.v[2], one.v[3]);
    if (lane == 1) one_w = make_uint4(one.v[4], one.v[5], one.v[6], one.v[7]);
    uint4 v = make_uint4(0, 0, 0, 0);
    if (good) { if (lane < 24u) v = q_lds128(wbase + 64u * BN_COOP_SLOT_F + 16u * lane); }
    else if (mode != 2) v = one_w;
    if (out && lane < 24u) reinterpret_cast<uint4*>(out + i)[lane] = v;
    const bool same = lane >= 24u || (v.x == one_w.x && v.y == one_w.y && v.z == one_w.z && v.w == one_w.w);
    const bool all_same = __all_sync(0xffffffffu, same);
    if (is_one && lane == 0) is_one[i] = all_same ? 1 : 0;
    if (ok && lane == 0) ok[i] = good ? 1 : 0;
}

// Gt::pow on a group (k_coop_gt_pow of bn_coop.cuh, same walk over the canonical scalar)
__global__ void __launch_bounds__(BN_COOP4_MAX_GROUPS * BN_COOP4_GROUP, 1) k_coop4_gt_pow(const Fq12* __restrict__ a, const Fq* __restrict__ fr,
                                                                                         Fq12* __restrict__ out, size_t n) {
    const sp tab = coop4_load_tables();
    const sp s_code = tab, s_hdr = tab + 4u * BN_COOP4_CODE_WORDS;
    const unsigned group = threadIdx.x / BN_COOP4_GROUP, gt = threadIdx.x % BN_COOP4_GROUP;
    const unsigned vlane = ((((gt >> 5) + group) & 3u) << 3) | ((gt >> 2) & 7u), part = gt & 3u;   // lanes 0..7 (the busiest) of the block's groups on different sub-partitions
    const size_t i = (size_t)blockIdx.x * (blockDim.x / BN_COOP4_GROUP) + group;
    if (i >= n) return;
    const sp wbase = tab + 4u * BN_COOP4_TABLE_WORDS + group * BN_COOP_SLOT_BYTES;
    for (unsigned k = gt; k < BN_COOP_NCONST * 16u; k += BN_COOP4_GROUP) q_sts32(wbase + 4u * k, bn_coop_const[k]);
    const uint32_t* src = reinterpret_cast<const uint32_t*>(a + i);
    for (unsigned k = gt; k < 96u; k += BN_COOP4_GROUP) q_sts32(wbase + 64u * BN_COOP_SLOT_EREG1 + 4u * k, src[k]);
    Fq m = fr[i];
    uint32_t e[8];
    fr_to_canonical(e, m.v);                                     // every thread the same scalar: the walk below is uniform
    {
        const Fq one = fq_one();
        uint4 v = make_uint4(0, 0, 0, 0);
        if (gt == 0) v = make_uint4(one.v[0], one.v[1], one.v[2], one.v[3]);
        if (gt == 1) v = make_uint4(one.v[4], one.v[5], one.v[6], one.v[7]);
        if (gt < 24u) q_sts128(wbase + 64u * BN_COOP_SLOT_F + 16u * gt, v);
    }
    coop4_sync(group);
    bool started = false;
    const uint32_t mul_bases = ((uint32_t)BN_COOP_SLOT_F << 8) | ((uint32_t)BN_COOP_SLOT_F << 16) | ((uint32_t)BN_COOP_SLOT_EREG1 << 24);
    // SQR12B takes a0 + v a1 of the accumulator from the slot T0, which SQR12B, MUL12T and T0INIT leave behind: one LIN step less
#pragma unroll 1
    for (int b = 255; b >= 0; b--) {
        if (started) coop4_call(BN_COOP_R_SQR12B_BEGIN, BN_COOP_R_SQR12B_COUNT, 0u, s_code, s_hdr, wbase, vlane, part, group);
        if ((e[b >> 5] >> (b & 31)) & 1u) {
            if (started) {
                coop4_call(BN_COOP_R_MUL12T_BEGIN, BN_COOP_R_MUL12T_COUNT, mul_bases, s_code, s_hdr, wbase, vlane, part, group);
            } else {
                if (gt < 24u) q_sts128(wbase + 64u * BN_COOP_SLOT_F + 16u * gt, q_lds128(wbase + 64u * BN_COOP_SLOT_EREG1 + 16u * gt));
                coop4_sync(group);
                coop4_call(BN_COOP_R_T0INIT_BEGIN, BN_COOP_R_T0INIT_COUNT, 0u, s_code, s_hdr, wbase, vlane, part, group);
                started = true;
            }
        }
    }
    if (gt < 24u) reinterpret_cast<uint4*>(out + i)[gt] = q_lds128(wbase + 64u * BN_COOP_SLOT_F + 16u * gt);
}

}  // namespace bn
#endif
