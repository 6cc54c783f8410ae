// Warp-cooperative pairing: the latency path for small batches (BASELINE configs[0], src/lib.rs:611-613 at n = 1;
// examples/joux.rs:17-21).
//
// The throughput kernels give one THREAD a whole pairing: a 17,730-multiplication dependent chain, ~10 ms however few
// pairings there are.  Here one WARP computes one pairing as a 32-lane vector machine over Fq2: a per-warp file of
// BN_COOP_NSLOTS Fq2 slots in shared memory, and two instruction kinds executed in lock step, every lane with its own operands
//     MUL  dst = conjA?(sum +-slot[a_t]) * conjB?(sum +-slot[b_t])        one Fq2 product per lane
//     LIN  dst = half?(P + xi X)                                          P, X = signed sums of slots
// plus INV (one lane, binary extended Euclid in registers).  The 18 Fq2 products of an Fq12 multiplication, the 12 of a squaring,
// the 9 of a cyclotomic squaring, the 18 of a line multiplication and the products of a G2 line step run on as many lanes
// at once; the dependent chain shrinks from 17,730 Fq multiplications to 1,651 steps.
//
// This file is ONLY the executor: it knows nothing of pairings.  The schedule (which lane multiplies what, in which slot)
// is the generated table bn_coop_ucode.cuh; tools/gen_coop_ucode.py holds the algorithm (the reference's formulas, cited
// there), simulates the very same tables on Python integers against the oracle (tests/test_coop_ucode.py) and emits them.
// Same field elements as the throughput path, hence the same bytes.
#pragma once
#include "bn_coop_ucode.cuh"
#include "bn_slots_curve.cuh"

#if defined(__CUDACC__)
namespace bn {

#define BN_HAVE_COOP 1
#define BN_COOP_MAX_WARPS 16                                   /* 16 x 11,136 B of slots + 46 KB of tables = 224 KB: one block per SM */
#define BN_COOP_SLOT_BYTES (BN_COOP_NSLOTS * 64)
#define BN_COOP_CODE_WORDS (BN_COOP_NSTEPS_SEQ * 160)            /* the steps of this executor: a prefix of the table (the rest is bn_coop4.cuh's) */
#define BN_COOP_TABLE_WORDS ((BN_COOP_CODE_WORDS + BN_COOP_NSTEPS_SEQ + BN_COOP_NTRACE + 3) & ~3)
#define BN_COOP_SMEM(warps) ((size_t)BN_COOP_TABLE_WORDS * 4 + (size_t)(warps) * BN_COOP_SLOT_BYTES)
#define BN_COOP_DEFAULT_MAX_N 2368                             /* up to one warp per pairing on every SM (148 x 16) */

// Accessors of the cooperative kernels.  Reads may touch the block's read-only tables or the warp's own slot file, writes only
// the warp's own file (checked in -DBN_SLOT_CHECK builds, bn_slots.cuh).
#ifdef BN_SLOT_CHECK
__device__ __forceinline__ bool coop_ok(sp a, unsigned bytes, bool write) {
    const sp dyn0 = (sp)__cvta_generic_to_shared(bn_slot_mem);
    const sp file = dyn0 + 4u * BN_COOP_TABLE_WORDS + (threadIdx.x >> 5) * BN_COOP_SLOT_BYTES;
    if (!write && a < dyn0 + 4u * BN_COOP_TABLE_WORDS) return slot_ok(3u, a, bytes, dyn0, 4u * BN_COOP_TABLE_WORDS);
    return slot_ok(write ? 5u : 4u, a, bytes, file, BN_COOP_SLOT_BYTES);
}
#endif
__device__ __forceinline__ uint32_t lds32(sp a) {
    uint32_t v = 0;
    BN_SLOT_GUARD(coop_ok(a, 4u, false)) return v;
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(a) : "memory");
    return v;
}
__device__ __forceinline__ void sts32(sp a, uint32_t v) {
    BN_SLOT_GUARD(coop_ok(a, 4u, true)) return;
    asm volatile("st.shared.u32 [%0], %1;" :: "r"(a), "r"(v) : "memory");
}
__device__ __forceinline__ uint4 c_lds128(sp a) {
    BN_SLOT_GUARD(coop_ok(a, 16u, false)) return make_uint4(0, 0, 0, 0);
    return lds128_raw(a);
}
__device__ __forceinline__ void c_sts128(sp a, uint4 v) {
    BN_SLOT_GUARD(coop_ok(a, 16u, true)) return;
    sts128_raw(a, v);
}
__device__ __forceinline__ Fq2 c_ld(sp a) {
    uint4 w0 = c_lds128(a), w1 = c_lds128(a + 16), w2 = c_lds128(a + 32), w3 = c_lds128(a + 48);
    Fq2 r;
    r.c0.v[0] = w0.x; r.c0.v[1] = w0.y; r.c0.v[2] = w0.z; r.c0.v[3] = w0.w;
    r.c0.v[4] = w1.x; r.c0.v[5] = w1.y; r.c0.v[6] = w1.z; r.c0.v[7] = w1.w;
    r.c1.v[0] = w2.x; r.c1.v[1] = w2.y; r.c1.v[2] = w2.z; r.c1.v[3] = w2.w;
    r.c1.v[4] = w3.x; r.c1.v[5] = w3.y; r.c1.v[6] = w3.z; r.c1.v[7] = w3.w;
    return r;
}
__device__ __forceinline__ void c_st(sp a, const Fq2& v) {
    c_sts128(a, make_uint4(v.c0.v[0], v.c0.v[1], v.c0.v[2], v.c0.v[3]));
    c_sts128(a + 16, make_uint4(v.c0.v[4], v.c0.v[5], v.c0.v[6], v.c0.v[7]));
    c_sts128(a + 32, make_uint4(v.c1.v[0], v.c1.v[1], v.c1.v[2], v.c1.v[3]));
    c_sts128(a + 48, make_uint4(v.c1.v[4], v.c1.v[5], v.c1.v[6], v.c1.v[7]));
}
// acc +- slot named by the 16-bit term t (slot code | minus << 10); bases: byte k = base of selector k (byte 0 = 0)
__device__ __forceinline__ void coop_acc(Fq2& acc, uint32_t t, sp wbase, uint32_t bases) {
    const uint32_t slot = (t & 0xffu) + ((bases >> (8u * ((t >> 8) & 3u))) & 0xffu);
    const Fq2 x = c_ld(wbase + slot * 64u);
    if (t & 0x400u) { fq_sub(acc.c0, acc.c0, x.c0); fq_sub(acc.c1, acc.c1, x.c1); }
    else { fq_add(acc.c0, acc.c0, x.c0); fq_add(acc.c1, acc.c1, x.c1); }
}

// One step of the microcode for this lane: decode from the shared-memory tables, accumulate the operand sums, compute, store.
__device__ __forceinline__ void coop_step(uint32_t tr, sp s_code, sp s_hdr, sp wbase, unsigned lane) {
    const uint32_t step = tr & 0xffu, bases = tr & 0xffffff00u;   // step | base D << 8 | base A << 16 | base B << 24
    const uint32_t hdr = lds32(s_hdr + 4u * step);               // op | n0 << 4 | n1 << 8 | any conj << 12 | any half << 13
    const sp cw = s_code + 4u * (step * 160u + lane);
    const uint32_t w0 = lds32(cw);
    unsigned long long tq_lo = (unsigned long long)lds32(cw + 128u) | ((unsigned long long)lds32(cw + 256u) << 32);
    unsigned long long tq_hi = (unsigned long long)lds32(cw + 384u) | ((unsigned long long)lds32(cw + 512u) << 32);
    const unsigned op = hdr & 15u, n0 = (hdr >> 4) & 15u, n1 = (hdr >> 8) & 15u;
    Fq2 x, y, r;
    x.c0 = fq_zero(); x.c1 = fq_zero(); y = x;
#pragma unroll 1
    for (unsigned k = 0; k < n0; k++) {
        coop_acc(x, (uint32_t)tq_lo & 0xffffu, wbase, bases);
        tq_lo = (tq_lo >> 16) | (tq_hi << 48); tq_hi >>= 16;
    }
#pragma unroll 1
    for (unsigned k = 0; k < n1; k++) {
        coop_acc(y, (uint32_t)tq_lo & 0xffffu, wbase, bases);
        tq_lo = (tq_lo >> 16) | (tq_hi << 48); tq_hi >>= 16;
    }
    const bool active = (w0 >> 10) & 1u;
    if (op == 1u) {                                              // MUL
        if (hdr & 0x1000u) {
            Fq m;
            fq_neg(m, x.c1); if (w0 & 0x800u) x.c1 = m;
            fq_neg(m, y.c1); if (w0 & 0x1000u) y.c1 = m;
        }
        fq2_mul_inl(r, x, y);
    } else if (op == 2u) {                                       // LIN
        if (n1) {
            Fq2 z;
            fq_lin9<true>(z.c0, y.c0, y.c1);
            fq_lin9<false>(z.c1, y.c1, y.c0);
            fq_add(r.c0, x.c0, z.c0); fq_add(r.c1, x.c1, z.c1);
        } else {
            r = x;
        }
        if (hdr & 0x2000u) {
            Fq2 h;
            fq_half_r(h.c0, r.c0); fq_half_r(h.c1, r.c1);
            if (w0 & 0x2000u) r = h;
        }
    } else {                                                     // INV: one lane
        r.c0 = fq_zero(); r.c1 = fq_zero();
        if (active) fq_inv_bgcd(r.c0, x.c0);
    }
    if (active) {
        const uint32_t slot = (w0 & 0xffu) + ((bases >> (8u * ((w0 >> 8) & 3u))) & 0xffu);
        c_st(wbase + slot * 64u, r);
    }
}

// the tables travel to shared memory once per block (46 KB): every step then decodes from on-chip memory
__device__ __forceinline__ sp coop_load_tables() {
    uint32_t* dst = reinterpret_cast<uint32_t*>(bn_slot_mem);
    for (unsigned k = threadIdx.x; k < BN_COOP_CODE_WORDS; k += blockDim.x) dst[k] = bn_coop_code[k];
    for (unsigned k = threadIdx.x; k < BN_COOP_NSTEPS_SEQ; k += blockDim.x) dst[BN_COOP_CODE_WORDS + k] = bn_coop_hdr[k];
    for (unsigned k = threadIdx.x; k < BN_COOP_NTRACE; k += blockDim.x) dst[BN_COOP_CODE_WORDS + BN_COOP_NSTEPS_SEQ + k] = bn_coop_trace[k];
    __syncthreads();
    return (sp)__cvta_generic_to_shared(bn_slot_mem);
}
// run the `count` steps of one routine starting at step `begin` with the bases (D, A, B) packed as D << 8 | A << 16 | B << 24
__device__ __forceinline__ void coop_call(unsigned begin, unsigned count, uint32_t bases, sp s_code, sp s_hdr, sp wbase, unsigned lane) {
#pragma unroll 1
    for (unsigned k = 0; k < count; k++) {
        coop_step((begin + k) | bases, s_code, s_hdr, wbase, lane);
        __syncwarp();
    }
}

// mode 0: out[i] = pairing(g1[i], g2[i]) (one for a pair with a point at infinity, src/groups/mod.rs:896)
// mode 1: out[i] = the un-exponentiated Miller value of the pair (one, ok[i] = 0 at infinity)
// mode 2: out[i] = final_exponentiation(in[i]), in passed through g1_or_gt (may alias out); ok[i] = 0 and zero for in == 0 (None)
// out, ok, is_one may each be null; is_one[i] = (result == Gt::one()).
__global__ void __launch_bounds__(BN_COOP_MAX_WARPS * 32, 1) k_coop_pairing(const void* g1_or_gt, const G2* __restrict__ g2, Fq12* out,
                                                                            uint8_t* __restrict__ ok, uint8_t* __restrict__ is_one, size_t n, int mode) {
    const sp tab = coop_load_tables();
    const sp s_code = tab, s_hdr = tab + 4u * BN_COOP_CODE_WORDS, s_trace = s_hdr + 4u * BN_COOP_NSTEPS_SEQ;
    const unsigned lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    const size_t i = (size_t)blockIdx.x * (blockDim.x >> 5) + warp;
    if (i >= n) return;                                          // whole warps leave: nothing below synchronises the block
    const sp wbase = tab + 4u * BN_COOP_TABLE_WORDS + warp * BN_COOP_SLOT_BYTES;

    // constants and inputs into the slot file (16 words per slot)
    for (unsigned k = lane; k < BN_COOP_NCONST * 16u; k += 32u) sts32(wbase + 4u * k, bn_coop_const[k]);
    bool good = true;
    int t_lo, t_hi;
    if (mode == 2) {
        const uint32_t* src = reinterpret_cast<const uint32_t*>(reinterpret_cast<const Fq12*>(g1_or_gt) + i);
        uint32_t any = 0;
        for (unsigned k = lane; k < 96u; k += 32u) {
            const uint32_t v = src[k];
            any |= v;
            sts32(wbase + 64u * BN_COOP_SLOT_F + 4u * k, v);
        }
        good = __ballot_sync(0xffffffffu, any != 0) != 0;        // zero has no final exponentiation (Option::None)
        t_lo = BN_COOP_T_FEXP_BEGIN; t_hi = BN_COOP_T_FEXP_END;
    } else {
        const uint32_t* p = reinterpret_cast<const uint32_t*>(reinterpret_cast<const G1*>(g1_or_gt) + i);   // x, y, z: 8 words each
        const uint32_t* q = reinterpret_cast<const uint32_t*>(g2 + i);                                      // x, y, z: 16 words each
        // G1 coordinates become (v, 0): 3 slots x 16 words, the upper 8 zero
        for (unsigned k = lane; k < 48u; k += 32u) {
            const unsigned c = k >> 4, w = k & 15u;
            sts32(wbase + 64u * (BN_COOP_SLOT_IN_P + c) + 4u * w, w < 8u ? p[8u * c + w] : 0u);
        }
        for (unsigned k = lane; k < 48u; k += 32u) sts32(wbase + 64u * BN_COOP_SLOT_IN_Q + 4u * k, q[k]);
        // a point at infinity (z = 0) on either side: the result is one (every lane evaluates the same words)
        uint32_t pz = 0, qz = 0;
        for (unsigned w = 0; w < 8u; w++) pz |= p[16u + w];
        for (unsigned w = 0; w < 16u; w++) qz |= q[32u + w];
        good = pz != 0 && qz != 0;
        t_lo = BN_COOP_T_PAIRING_BEGIN;
        t_hi = mode == 1 ? BN_COOP_T_MILLER_END : BN_COOP_T_PAIRING_END;
    }
    __syncwarp();

    if (good) {
#pragma unroll 1
        for (int t = t_lo; t < t_hi; t++) {
            coop_step(lds32(s_trace + 4u * (unsigned)t), s_code, s_hdr, wbase, lane);
            __syncwarp();                                        // the one synchronisation of the step
        }
    }
    // word `lane` of Gt::one(): (1, 0) in the first slot, zero elsewhere
    const Fq one = fq_one();
    uint4 one_w = make_uint4(0, 0, 0, 0);
    if (lane == 0) one_w = make_uint4(one.v[0], one.v[1], one.v[2], one.v[3]);
    if (lane == 1) one_w = make_uint4(one.v[4], one.v[5], one.v[6], one.v[7]);
    uint4 v = make_uint4(0, 0, 0, 0);
    if (good) { if (lane < 24u) v = c_lds128(wbase + 64u * BN_COOP_SLOT_F + 16u * lane); }
    else if (mode != 2) v = one_w;                               // infinity -> one; a zero input of mode 2 -> zero
    if (out && lane < 24u) reinterpret_cast<uint4*>(out + i)[lane] = v;          // 384 contiguous bytes per warp
    const bool same = lane >= 24u || (v.x == one_w.x && v.y == one_w.y && v.z == one_w.z && v.w == one_w.w);
    const bool all_same = __all_sync(0xffffffffu, same);
    if (is_one && lane == 0) is_one[i] = all_same ? 1 : 0;
    if (ok && lane == 0) ok[i] = good ? 1 : 0;
}

// Gt::pow (src/lib.rs:592 -> FieldElement::pow, src/fields/mod.rs:35-46) one warp per element: MSB-first square-and-multiply over
// the canonical scalar with the SQR12 / MUL12 routines of the microcode (accumulator in the F slots, base in register 1);
// leading zero bits only square one, so skipping them gives the same value (as fq12_pow of bn_tower.cuh does).
__global__ void __launch_bounds__(BN_COOP_MAX_WARPS * 32, 1) k_coop_gt_pow(const Fq12* __restrict__ a, const Fq* __restrict__ fr,
                                                                           Fq12* __restrict__ out, size_t n) {
    const sp tab = coop_load_tables();
    const sp s_code = tab, s_hdr = tab + 4u * BN_COOP_CODE_WORDS;
    const unsigned lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    const size_t i = (size_t)blockIdx.x * (blockDim.x >> 5) + warp;
    if (i >= n) return;
    const sp wbase = tab + 4u * BN_COOP_TABLE_WORDS + warp * BN_COOP_SLOT_BYTES;
    for (unsigned k = lane; k < BN_COOP_NCONST * 16u; k += 32u) sts32(wbase + 4u * k, bn_coop_const[k]);
    const uint32_t* src = reinterpret_cast<const uint32_t*>(a + i);
    for (unsigned k = lane; k < 96u; k += 32u) sts32(wbase + 64u * BN_COOP_SLOT_EREG1 + 4u * k, src[k]);
    Fq m = fr[i];
    uint32_t e[8];
    fr_to_canonical(e, m.v);                                     // every lane the same scalar: the walk below is uniform
    // F = one
    {
        const Fq one = fq_one();
        uint4 v = make_uint4(0, 0, 0, 0);
        if (lane == 0) v = make_uint4(one.v[0], one.v[1], one.v[2], one.v[3]);
        if (lane == 1) v = make_uint4(one.v[4], one.v[5], one.v[6], one.v[7]);
        if (lane < 24u) c_sts128(wbase + 64u * BN_COOP_SLOT_F + 16u * lane, v);
    }
    __syncwarp();
    bool started = false;
    const uint32_t mul_bases = ((uint32_t)BN_COOP_SLOT_F << 8) | ((uint32_t)BN_COOP_SLOT_F << 16) | ((uint32_t)BN_COOP_SLOT_EREG1 << 24);
#pragma unroll 1
    for (int b = 255; b >= 0; b--) {
        if (started) coop_call(BN_COOP_R_SQR12_BEGIN, BN_COOP_R_SQR12_COUNT, 0u, s_code, s_hdr, wbase, lane);
        if ((e[b >> 5] >> (b & 31)) & 1u) {
            if (started) {
                coop_call(BN_COOP_R_MUL12_BEGIN, BN_COOP_R_MUL12_COUNT, mul_bases, s_code, s_hdr, wbase, lane);
            } else {
                if (lane < 24u) c_sts128(wbase + 64u * BN_COOP_SLOT_F + 16u * lane, c_lds128(wbase + 64u * BN_COOP_SLOT_EREG1 + 16u * lane));
                __syncwarp();
                started = true;
            }
        }
    }
    if (lane < 24u) reinterpret_cast<uint4*>(out + i)[lane] = c_lds128(wbase + 64u * BN_COOP_SLOT_F + 16u * lane);
}

}  // namespace bn
#endif
