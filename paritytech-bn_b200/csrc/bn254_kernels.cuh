// CUDA kernels of the BN254 pairing engine (sm_100a).  One thread per unit of work (pairing,
// product check, scalar multiplication); Fq12 state in per-thread local memory, Fq2-level bodies
// register-blocked around the IMAD.WIDE carry chains of fq_ptx.cuh.  See DESIGN.md for the layout.
#pragma once
#include "bn_curve.cuh"
#include "bn_wire.cuh"
#include "bn_slots.cuh"
#include "bn_slots_curve.cuh"
#include "bn_coop4.cuh"

namespace bn {

#ifndef BN_BLOCK
#define BN_BLOCK 64
#endif
#ifndef BN_MIN_BLOCKS
#define BN_MIN_BLOCKS 8
#endif
#define BN_KERNEL __global__ void __launch_bounds__(BN_BLOCK, BN_MIN_BLOCKS)
// the final exponentiation has no state besides Fq12 values in local memory: 96 registers (10 blocks/SM) cost no
// spills and give 5 warps per sub-partition on batches > 2^16 (+5 %, profiles/r01_bench_history.md)
#ifndef BN_MIN_BLOCKS_FEXP
#define BN_MIN_BLOCKS_FEXP 10
#endif
#define BN_KERNEL_FEXP __global__ void __launch_bounds__(BN_BLOCK, BN_MIN_BLOCKS_FEXP)
#ifndef BN_MIN_BLOCKS_LINES
#define BN_MIN_BLOCKS_LINES BN_MIN_BLOCKS
#endif
#ifndef BN_MIN_BLOCKS_MTAB
#define BN_MIN_BLOCKS_MTAB BN_MIN_BLOCKS
#endif

__device__ __forceinline__ size_t gtid() { return (size_t)blockIdx.x * blockDim.x + threadIdx.x; }

// ---------------------------------------------------------------------------------------------
// affine conversion of one (G1, G2) pair with ONE Fermat inversion (the reference does two,
// src/groups/mod.rs:895; same field elements).  false if either point is at infinity.
// ---------------------------------------------------------------------------------------------
__device__ __noinline__ bool pair_to_affine(AffG1& pa, AffG2& qa, const G1& p, const G2& q) {
    if (fq_is_zero(p.z) || fq2_is_zero(q.z)) return false;
    Fq n, t, ti, pzi, ni, z2, z3;
    fq_mul(n, q.z.c0, q.z.c0);
    fq_mul(t, q.z.c1, q.z.c1);
    fq_add(n, n, t);                 // norm(q.z) != 0
    fq_mul_ni(t, p.z, n);
    fq_inv_bgcd(ti, t);               // binary extended Euclid in 17 rounds of 30 approximated steps: same element
    fq_mul_ni(pzi, ti, n);           // 1/p.z
    fq_mul_ni(ni, ti, p.z);          // 1/norm
    fq_mul_ni(z2, pzi, pzi);
    fq_mul_ni(pa.x, p.x, z2);
    fq_mul_ni(z3, z2, pzi);
    fq_mul_ni(pa.y, p.y, z3);
    Fq2 qzi, w2, w3;
    fq_mul_ni(qzi.c0, q.z.c0, ni);
    fq_mul_ni(t, q.z.c1, ni);
    fq_neg(qzi.c1, t);               // 1/q.z = conj(q.z)/norm
    fq2_sqr(w2, qzi);
    fq2_mul(qa.x, q.x, w2);
    fq2_mul(w3, w2, qzi);
    fq2_mul(qa.y, q.y, w3);
    return true;
}

// ---------------------------------------------------------------------------------------------
// pairing = Miller kernel + final-exponentiation kernel (the 384-byte Miller value makes one trip
// through HBM: negligible against ~4.7 M IMADs per pairing, and each phase gets its own register
// allocation).  lib.rs:611-613 / groups/mod.rs:894-902.
// ---------------------------------------------------------------------------------------------
BN_KERNEL k_miller_many(const G2* __restrict__ q, const G1* __restrict__ p, Fq12* __restrict__ f_out,
                        uint8_t* __restrict__ ok, size_t n) {
    size_t i = gtid();
    if (i >= n) return;
    G1 pp = p[i];
    G2 qq = q[i];
    AffG1 pa;
    AffG2 qa;
    Fq12 f = fq12_one();
    bool good = pair_to_affine(pa, qa, pp, qq);
    if (good) miller_loop_fused(f, pa, qa);
    f_out[i] = f;
    if (ok) ok[i] = good ? 1 : 0;
}

BN_KERNEL_FEXP k_final_exp_many(const Fq12* __restrict__ in, Fq12* __restrict__ out, uint8_t* __restrict__ some,
                           uint8_t* __restrict__ is_one, size_t n) {
    size_t i = gtid();
    if (i >= n) return;
    Fq12 f = in[i], r;
    bool s = fq12_final_exponentiation(r, f);
    if (out) out[i] = r;
    if (some) some[i] = s ? 1 : 0;
    if (is_one) {
        Fq12 one = fq12_one();
        is_one[i] = fq12_eq(r, one) ? 1 : 0;
    }
}

// final exponentiation as two kernels: the easy part carries the inversion / Frobenius code (cold), the hard
// part is the x^u loops; keeping the cold code out of the hard-part kernel keeps its hot SASS contiguous.
BN_KERNEL k_final_exp_easy(const Fq12* __restrict__ in, Fq12* __restrict__ out, uint8_t* __restrict__ some, size_t n) {
    size_t i = gtid();
    if (i >= n) return;
    Fq12 f = in[i], r;
    bool s = fq12_final_exp_first_chunk(r, f);
    out[i] = r;
    some[i] = s ? 1 : 0;
}
BN_KERNEL_FEXP k_final_exp_hard(const Fq12* __restrict__ in, const uint8_t* __restrict__ some_in, Fq12* __restrict__ out,
                           uint8_t* __restrict__ some, uint8_t* __restrict__ is_one, size_t n) {
    size_t i = gtid();
    if (i >= n) return;
    bool s = some_in[i] != 0;
    Fq12 f = in[i], r;
    if (s) fq12_final_exp_last_chunk(r, f); else r = fq12_zero();
    if (out) out[i] = r;
    if (some) some[i] = s ? 1 : 0;
    if (is_one) {
        Fq12 one = fq12_one();
        is_one[i] = fq12_eq(r, one) ? 1 : 0;
    }
}

// hard part on the shared-memory slot file (bn_slots.cuh): launched with BN_BLOCK * BN_SLOT_STRIDE bytes of dynamic
// shared memory, 7 blocks per SM
__global__ void __launch_bounds__(BN_BLOCK, 7) k_final_exp_hard_s(const Fq12* __restrict__ in, const uint8_t* __restrict__ some_in,
                           Fq12* __restrict__ out, uint8_t* __restrict__ some, uint8_t* __restrict__ is_one, size_t n) {
    size_t i = gtid();
    if (i >= n) return;
    bool s = some_in[i] != 0;
    Fq12 f = in[i], r;
    if (s) s_fq12_final_exp_last_chunk<false>(r, f, s_base()); else r = fq12_zero();
    if (out) out[i] = r;
    if (some) some[i] = s ? 1 : 0;
    if (is_one) {
        Fq12 one = fq12_one();
        is_one[i] = fq12_eq(r, one) ? 1 : 0;
    }
}

// same, one block of BN_FEXP_BLOCK threads per SM with phase barriers (bn_slots.cuh, s_phase): threads past n run the
// schedule on zeros so that every barrier is reached by the whole block
#ifndef BN_FEXP_BLOCK
#define BN_FEXP_BLOCK 448               /* one-wave block: 7 slots per thread */
#endif
#define BN_SYNC_MAX_BLOCK 512           /* batches of more than one wave: 16 warps per SM on 6 slots per thread */
__global__ void __launch_bounds__(BN_SYNC_MAX_BLOCK, 1) k_final_exp_hard_sync(const Fq12* __restrict__ in, const uint8_t* __restrict__ some_in,
                           Fq12* __restrict__ out, uint8_t* __restrict__ some, uint8_t* __restrict__ is_one, size_t n) {
    size_t i = gtid();
    bool live = i < n;
    bool s = live && some_in[i] != 0;
    Fq12 f, r;
    if (s) f = in[i]; else f = fq12_zero();
    const sp base = s_base();
    s_fq12_final_exp_last_chunk<true>(r, f, base);
    // Results leave through the slot file: every thread parks its 384 bytes in its own slots, then each warp stores its 32
    // consecutive records (12 KB, contiguous in the AoS output) as 24 fully coalesced 512-byte rows.  A per-thread store is
    // 16 bytes at a 384-byte stride per lane -- half-filled sectors in HBM and, for a pinned (zero-copy) output buffer,
    // 16-byte PCIe writes.
    if (out) {
        if (!s) r = fq12_zero();
        const uint4* rw = reinterpret_cast<const uint4*>(&r);
#pragma unroll
        for (int w = 0; w < 24; w++) sts128(base + 16 * w, rw[w]);
        __syncwarp();
        const unsigned stride = blockDim.x > BN_SLOT_BLOCK7 ? BN_SLOT_STRIDE6 : BN_SLOT_STRIDE;
        const unsigned lane = threadIdx.x & 31u;
        const sp wbase = base - lane * stride;                       // slots of lane 0 of this warp
        const size_t first = i - lane;                               // record index of lane 0
        uint4* o = reinterpret_cast<uint4*>(out + first);
#pragma unroll 4
        for (int row = 0; row < 24; row++) {
            unsigned g = row * 32u + lane;                           // 16-byte word of the warp's 12 KB
            unsigned t = g / 24u, w = g - t * 24u;
            if (first + t < n) {
                const sp src = wbase + t * stride + 16 * w;          // another lane's slots: the warp's 32 regions
                BN_SLOT_GUARD(slot_ok(2u, src, 16u, wbase, 32u * stride)) continue;
                o[g] = lds128_raw(src);
            }
        }
    }
    if (!live) return;
    if (some) some[i] = s ? 1 : 0;
    if (is_one) {
        Fq12 one = fq12_one();
        is_one[i] = fq12_eq(r, one) ? 1 : 0;
    }
}

// Miller loop from the line table on the slot file, one phase-synchronised block per SM: the squaring code and the
// line-multiplication code alternate under block barriers so that each fits the instruction cache on its own
// (G2Precomp::miller_loop, src/groups/mod.rs:580-606; same values as k_miller_table_affine).
__global__ void __launch_bounds__(BN_SYNC_MAX_BLOCK, 1) k_miller_table_sync(const AffG1* __restrict__ pa_in, const uint8_t* __restrict__ ok_in,
                                const uint4* __restrict__ tab, Fq12* __restrict__ f_out, uint8_t* __restrict__ ok_out, size_t n) {
    size_t i = gtid();
    bool live = i < n;
    size_t j = live ? i : n - 1;
    bool good = live && ok_in[j] != 0;
    AffG1 pa = pa_in[j];
    Fq12 f0 = fq12_one(), f1;
    bool in0 = true;
    const sp base = s_base();
    int idx = 0;
#define BN_LINE_STEP() do { const uint4* c = tab + (size_t)idx * 12 * n + j; idx++; \
        if (in0) s_line_mul(f1, f0, c, n, pa, base); else s_line_mul(f0, f1, c, n, pa, base); in0 = !in0; } while (0)
#pragma unroll 1
    for (int k = 0; k < 64; k++) {
        if (k) {
            __syncthreads();
            if (in0) s_fq12_sqr_na(f1, f0, base); else s_fq12_sqr_na(f0, f1, base);
            in0 = !in0;
            __syncthreads();
        }
        BN_LINE_STEP();
        if (naf_digit(k)) BN_LINE_STEP();
    }
    BN_LINE_STEP();
    BN_LINE_STEP();
#undef BN_LINE_STEP
    if (!live) return;
    if (!good) f0 = fq12_one(); else if (!in0) f0 = f1;
    f_out[i] = f0;
    if (ok_out) ok_out[i] = good ? 1 : 0;
}

// ---------------------------------------------------------------------------------------------
// product checks: one thread per check, shared squaring chain (groups::miller_loop_batch,
// src/groups/mod.rs:609-640), line coefficients of the variable G2 points produced on the fly,
// those of shared G2 points read from device tables (broadcast loads).
// Table layout: uint4 tab[(c * 12 + w) * n_tables + j] = 16-byte word w of coefficient c of table j
// (word-major SoA: coalesced when thread i reads table i, a broadcast when all read table j).
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ void load_coeff(EllCoeffs& c, const uint4* __restrict__ tab, size_t n_tables, size_t j, int idx) {
    uint4* dst = reinterpret_cast<uint4*>(&c);
#pragma unroll
    for (int w = 0; w < 12; w++) dst[w] = __ldg(&tab[((size_t)idx * 12 + w) * n_tables + j]);
}

template <int MAXV, int MAXS>
BN_KERNEL k_miller_product(const G1* __restrict__ g1v, const G2* __restrict__ g2v, int v,
                           const G1* __restrict__ g1s, const uint4* __restrict__ tab, int m,
                           Fq12* __restrict__ f_out, size_t n_checks) {
    size_t i = gtid();
    if (i >= n_checks) return;
    G2 R[MAXV];
    AffG2 Q[MAXV];
    AffG1 P[MAXV];
    AffG1 PS[MAXS];
    int js[MAXS];
    int nv = 0, ns = 0;
    for (int j = 0; j < v; j++) {       // skip semantics of pairing_batch, src/groups/mod.rs:909-920
        G1 pp = g1v[i * v + j];
        G2 qq = g2v[i * v + j];
        if (pair_to_affine(P[nv], Q[nv], pp, qq)) {
            R[nv].x = Q[nv].x; R[nv].y = Q[nv].y; R[nv].z = fq2_one();
            nv++;
        }
    }
    for (int j = 0; j < m; j++) {
        G1 pp = g1s[i * m + j];
        if (jac_to_affine(PS[ns], pp)) js[ns++] = j;
    }
    Fq12 f = fq12_one();
    EllCoeffs c;
    int idx = 0;
    bool started = false;
    for (int k = 0; k < 64; k++) {
        if (started) fq12_sqr(f, f);
        for (int j = 0; j < nv; j++) { g2_doubling_step(R[j], c); line_mul(f, c, P[j]); started = true; }
        for (int j = 0; j < ns; j++) { load_coeff(c, tab, m, js[j], idx); line_mul(f, c, PS[j]); started = true; }
        idx++;
        uint32_t d = naf_digit(k);
        if (d != 0) {
            for (int j = 0; j < nv; j++) {
                AffG2 qn = Q[j];
                if (d == 3) fq2_neg(qn.y, qn.y);
                g2_mixed_addition_step(R[j], qn, c);
                line_mul(f, c, P[j]);
            }
            for (int j = 0; j < ns; j++) { load_coeff(c, tab, m, js[j], idx); line_mul(f, c, PS[j]); }
            idx++;
        }
    }
    for (int step = 0; step < 2; step++) {
        for (int j = 0; j < nv; j++) {
            AffG2 q1, q2;
            g2_mul_by_q(q1, Q[j]);
            if (step == 1) { g2_mul_by_q(q2, q1); fq2_neg(q2.y, q2.y); q1 = q2; }
            g2_mixed_addition_step(R[j], q1, c);
            line_mul(f, c, P[j]);
        }
        for (int j = 0; j < ns; j++) { load_coeff(c, tab, m, js[j], idx); line_mul(f, c, PS[j]); }
        idx++;
    }
    f_out[i] = f;
}

// ---------------------------------------------------------------------------------------------
// split Miller path (three kernels, each with < 32 KB of hot SASS so the loop bodies stay resident in
// the SM's 32 KB L1.5 instruction cache): (1) joint affine conversion, (2) line coefficients of Q_i
// into the word-major table (coalesced 16-byte stores), (3) f_i from table i and P_i.
// The table makes one trip through HBM: 16,704 B written + read per pairing (~2.2 GB per 2^16 batch).
// ---------------------------------------------------------------------------------------------
BN_KERNEL k_pairs_to_affine(const G2* __restrict__ q, const G1* __restrict__ p, AffG2* __restrict__ qa_out,
                            AffG1* __restrict__ pa_out, uint8_t* __restrict__ ok, size_t n) {
    size_t i = gtid();
    if (i >= n) return;
    G1 pp = p[i];
    G2 qq = q[i];
    AffG1 pa;
    AffG2 qa;
    bool good = pair_to_affine(pa, qa, pp, qq);
    if (good) { qa_out[i] = qa; pa_out[i] = pa; }
    ok[i] = good ? 1 : 0;
}
// pair i against table i (materialised-G2Precomp streaming variant: coalesced table reads)
BN_KERNEL k_miller_precomp_each(const G1* __restrict__ g1, const uint4* __restrict__ tab, size_t n_tables,
                                Fq12* __restrict__ f_out, uint8_t* __restrict__ ok, size_t n) {
    size_t i = gtid();
    if (i >= n) return;
    G1 pp = g1[i];
    AffG1 pa;
    Fq12 f = fq12_one();
    bool good = jac_to_affine(pa, pp);
    if (good) {
        EllCoeffs c;
        int idx = 0;
        for (int k = 0; k < 64; k++) {
            if (k) fq12_sqr(f, f);
            load_coeff(c, tab, n_tables, i, idx++); line_mul(f, c, pa);
            if (naf_digit(k)) { load_coeff(c, tab, n_tables, i, idx++); line_mul(f, c, pa); }
        }
        load_coeff(c, tab, n_tables, i, idx++); line_mul(f, c, pa);
        load_coeff(c, tab, n_tables, i, idx++); line_mul(f, c, pa);
    }
    f_out[i] = f;
    if (ok) ok[i] = good ? 1 : 0;
}

// AffineG2::precompute for n points into the word-major table; bad[0] set if any point is at infinity
struct TableSink {
    uint4* tab; size_t n_tables, j; int idx;
    __device__ __forceinline__ void put(const EllCoeffs& c) {
        const uint4* src = reinterpret_cast<const uint4*>(&c);
#pragma unroll
        for (int w = 0; w < 12; w++) tab[((size_t)idx * 12 + w) * n_tables + j] = src[w];
        idx++;
    }
    __device__ __forceinline__ void dbl(const EllCoeffs& c, int) { put(c); }
    __device__ __forceinline__ void add(const EllCoeffs& c) { put(c); }
};
BN_KERNEL k_g2_precompute(const G2* __restrict__ q, uint4* __restrict__ tab, AffG2* __restrict__ q_aff,
                          int* __restrict__ bad, size_t n) {
    size_t i = gtid();
    if (i >= n) return;
    G2 qq = q[i];
    AffG2 qa;
    if (!jac_to_affine(qa, qq)) { atomicExch(bad, 1); return; }
    q_aff[i] = qa;
    TableSink s;
    s.tab = tab; s.n_tables = n; s.j = i; s.idx = 0;
    g2_walk_lines(qa, s);
}

// Table slot of pair t: identity for independent pairings (v = 1); for product checks with v pairs per check
// (pair t = check t/v, position t%v) the slot is (t%v)*n_checks + t/v, so that the threads of the product
// kernel (one per check) read consecutive 16-byte words for every position.
__device__ __forceinline__ size_t table_slot(size_t t, size_t v, size_t n) { return v <= 1 ? t : (t % v) * (n / v) + t / v; }
__global__ void __launch_bounds__(BN_BLOCK, BN_MIN_BLOCKS_LINES) k_g2_lines_affine(const AffG2* __restrict__ qa, const uint8_t* __restrict__ ok,
                                                                                  uint4* __restrict__ tab, size_t n, size_t v) {
    size_t i = gtid();
    if (i >= n || !ok[i]) return;
    AffG2 q = qa[i];
    TableSink s;
    s.tab = tab; s.n_tables = n; s.j = table_slot(i, v, n); s.idx = 0;
    g2_walk_lines(q, s);
}
// Same walk on the slot file (bn_slots_curve.cuh), one phase-synchronised block per SM: doubling steps and addition steps
// alternate under block barriers (the NAF digits are constants, so every thread takes the same steps).  Threads without a
// valid point walk a dummy point and store nothing.
__global__ void __launch_bounds__(BN_FEXP_BLOCK, 1) k_g2_lines_sync(const AffG2* __restrict__ qa, const uint8_t* __restrict__ ok,
                                                                     uint4* __restrict__ tab, size_t n, size_t v) {
    size_t i = gtid();
    bool live = i < n && ok[i] != 0;
    AffG2 q, qn, q1, q2;
    if (live) q = qa[i]; else { q.x = fq2_one(); q.y = fq2_one(); }
    qn.x = q.x; fq2_neg(qn.y, q.y);
    g2_mul_by_q(q1, q);
    g2_mul_by_q(q2, q1);
    fq2_neg(q2.y, q2.y);
    const sp base = s_base();
    s_in(base, q.x); s_in(base + 64, q.y);
    { Fq2 one = fq2_one(); s_st(base + 128, one); }
    LineOut o;
    o.tab = tab + (live ? table_slot(i, v, n) : 0);
    o.stride = n;
    o.live = live;
    int idx = 0;
#pragma unroll 1
    for (int k = 0; k < 64; k++) {
        __syncthreads();
        s_g2_doubling_step(base, o, idx++);
        uint32_t d = naf_digit(k);
        if (d != 0) {
            __syncthreads();
            if (d == 1) s_g2_mixed_addition_step(base, q.x, q.y, o, idx++);
            else s_g2_mixed_addition_step(base, q.x, qn.y, o, idx++);
        }
    }
    __syncthreads();
    s_g2_mixed_addition_step(base, q1.x, q1.y, o, idx++);
    s_g2_mixed_addition_step(base, q2.x, q2.y, o, idx++);
}
// Product checks from line tables: one thread per check, shared squaring chain (groups::miller_loop_batch,
// src/groups/mod.rs:609-640).  Variable pairs: affine P / ok flags / per-pair tables from the two kernels above
// (slot layout of table_slot); shared pairs: Jacobian G1 points against the shared G2Precomp tables (broadcast loads).
template <int MAXV, int MAXS>
BN_KERNEL k_miller_product_tables(const AffG1* __restrict__ pa, const uint8_t* __restrict__ ok, const uint4* __restrict__ tabv, int v,
                                  const G1* __restrict__ g1s, const uint4* __restrict__ tabs, int m,
                                  Fq12* __restrict__ f_out, size_t n_checks) {
    size_t i = gtid();
    if (i >= n_checks) return;
    AffG1 P[MAXV];
    AffG1 PS[MAXS];
    int jv[MAXV], js[MAXS];
    int nv = 0, ns = 0;
    for (int j = 0; j < v; j++)          // skip semantics of pairing_batch, src/groups/mod.rs:909-920
        if (ok[i * v + j]) { P[nv] = pa[i * v + j]; jv[nv++] = j; }
    for (int j = 0; j < m; j++) {
        G1 pp = g1s[i * m + j];
        if (jac_to_affine(PS[ns], pp)) js[ns++] = j;
    }
    Fq12 f = fq12_one();
    EllCoeffs c;
    size_t nt = n_checks * (size_t)v;
    bool started = false;
#pragma unroll 1
    for (int idx = 0; idx < BN_N_COEFFS; idx++) {
        if (started && coeff_is_doubling(idx)) fq12_sqr(f, f);
        for (int j = 0; j < nv; j++) { load_coeff(c, tabv, nt, (size_t)jv[j] * n_checks + i, idx); line_mul(f, c, P[j]); started = true; }
        for (int j = 0; j < ns; j++) { load_coeff(c, tabs, m, js[j], idx); line_mul(f, c, PS[j]); started = true; }
    }
    f_out[i] = f;
}
// Same on the slot file (one phase-synchronised block per SM; see k_miller_table_sync below).  The barriers are uniform:
// every thread squares before every doubling coefficient but the first (a check without valid pairs squares 1) and runs
// its own number of line multiplications between the barriers.
template <int MAXV, int MAXS>
__global__ void __launch_bounds__(BN_SYNC_MAX_BLOCK, 1) k_miller_product_sync(const AffG1* __restrict__ pa, const uint8_t* __restrict__ ok,
                                  const uint4* __restrict__ tabv, int v, const G1* __restrict__ g1s, const uint4* __restrict__ tabs, int m,
                                  Fq12* __restrict__ f_out, size_t n_checks) {
    size_t i = gtid();
    bool live = i < n_checks;
    AffG1 P[MAXV];
    AffG1 PS[MAXS];
    int jv[MAXV], js[MAXS];
    int nv = 0, ns = 0;
    if (live) {
        for (int j = 0; j < v; j++)          // skip semantics of pairing_batch, src/groups/mod.rs:909-920
            if (ok[i * v + j]) { P[nv] = pa[i * v + j]; jv[nv++] = j; }
        for (int j = 0; j < m; j++) {
            G1 pp = g1s[i * m + j];
            if (jac_to_affine(PS[ns], pp)) js[ns++] = j;
        }
    }
    Fq12 f0 = fq12_one(), f1;
    bool in0 = true;
    const sp base = s_base();
    size_t nt = n_checks * (size_t)v;
#pragma unroll 1
    for (int idx = 0; idx < BN_N_COEFFS; idx++) {
        if (idx > 0 && coeff_is_doubling(idx)) {
            __syncthreads();
            if (in0) s_fq12_sqr_na(f1, f0, base); else s_fq12_sqr_na(f0, f1, base);
            in0 = !in0;
            __syncthreads();
        }
#pragma unroll 1
        for (int j = 0; j < nv; j++) {
            const uint4* c = tabv + (size_t)idx * 12 * nt + ((size_t)jv[j] * n_checks + i);
            if (in0) s_line_mul(f1, f0, c, nt, P[j], base); else s_line_mul(f0, f1, c, nt, P[j], base);
            in0 = !in0;
        }
#pragma unroll 1
        for (int j = 0; j < ns; j++) {
            const uint4* c = tabs + (size_t)idx * 12 * m + js[j];
            if (in0) s_line_mul(f1, f0, c, (size_t)m, PS[j], base); else s_line_mul(f0, f1, c, (size_t)m, PS[j], base);
            in0 = !in0;
        }
    }
    if (!live) return;
    if (!in0) f0 = f1;
    f_out[i] = f0;
}
__global__ void __launch_bounds__(BN_BLOCK, BN_MIN_BLOCKS_MTAB) k_miller_table_affine(const AffG1* __restrict__ pa_in, const uint8_t* __restrict__ ok_in, const uint4* __restrict__ tab,
                                Fq12* __restrict__ f_out, uint8_t* __restrict__ ok_out, size_t n) {
    size_t i = gtid();
    if (i >= n) return;
    Fq12 f = fq12_one();
    bool good = ok_in[i] != 0;
    if (good) {
        AffG1 pa = pa_in[i];
        EllCoeffs c;
        int idx = 0;
        for (int k = 0; k < 64; k++) {
            if (k) fq12_sqr(f, f);
            load_coeff(c, tab, n, i, idx++); line_mul(f, c, pa);
            if (naf_digit(k)) { load_coeff(c, tab, n, i, idx++); line_mul(f, c, pa); }
        }
        load_coeff(c, tab, n, i, idx++); line_mul(f, c, pa);
        load_coeff(c, tab, n, i, idx++); line_mul(f, c, pa);
    }
    f_out[i] = f;
    if (ok_out) ok_out[i] = good ? 1 : 0;
}

// out[i] = Fq12::one() (the empty product, src/groups/mod.rs:610, :922-924)
BN_KERNEL k_fq12_set_one(Fq12* __restrict__ out, size_t n) {
    size_t i = gtid();
    if (i >= n) return;
    out[i] = fq12_one();
}
// flag[0] = 1 if any of the n status bytes is zero (miller_loop_batch: a point at infinity anywhere -> Err, src/lib.rs:629-630)
__global__ void k_any_zero_flag(const uint8_t* __restrict__ ok, int* __restrict__ flag, size_t n) {
    size_t i = gtid();
    if (i < n && ok[i] == 0) atomicExch(flag, 1);
}
// out[g] = prod_{j < group} in[g*group + j]
BN_KERNEL k_fq12_group_product(const Fq12* __restrict__ in, size_t group, Fq12* __restrict__ out, size_t n_groups) {
    size_t g = gtid();
    if (g >= n_groups) return;
    Fq12 acc = in[g * group];
    for (size_t j = 1; j < group; j++) { Fq12 b = in[g * group + j]; fq12_mul(acc, acc, b); }
    out[g] = acc;
}
// one tree level: out[i] = in[2i] * in[2i+1] (last element copied when n is odd)
BN_KERNEL k_fq12_pair_product(const Fq12* __restrict__ in, size_t n, Fq12* __restrict__ out) {
    size_t i = gtid();
    size_t n_out = (n + 1) / 2;
    if (i >= n_out) return;
    Fq12 a = in[2 * i];
    if (2 * i + 1 < n) { Fq12 b = in[2 * i + 1]; fq12_mul(a, a, b); }
    out[i] = a;
}

// ---------------------------------------------------------------------------------------------
// Gt algebra (src/lib.rs:584-609), scalar multiplication (src/lib.rs:425-431, :576-582), normalize
// ---------------------------------------------------------------------------------------------
BN_KERNEL k_gt_mul(const Fq12* __restrict__ a, const Fq12* __restrict__ b, Fq12* __restrict__ out, size_t n) {
    size_t i = gtid();
    if (i >= n) return;
    Fq12 x = a[i], y = b[i];
    fq12_mul(x, x, y);
    out[i] = x;
}
BN_KERNEL k_gt_pow(const Fq12* __restrict__ a, const Fq* __restrict__ fr, Fq12* __restrict__ out, size_t n) {
    size_t i = gtid();
    if (i >= n) return;
    Fq12 x = a[i], r;
    Fq m = fr[i];
    uint32_t k[8];
    fr_to_canonical(k, m.v);
    fq12_pow(r, x, k);
    out[i] = r;
}
BN_KERNEL k_gt_inverse(const Fq12* __restrict__ a, Fq12* __restrict__ out, uint8_t* __restrict__ some, size_t n) {
    size_t i = gtid();
    if (i >= n) return;
    Fq12 x = a[i], r;
    bool s = fq12_inv(r, x);
    out[i] = r;
    if (some) some[i] = s ? 1 : 0;
}
__global__ void k_gt_eq(const Fq12* __restrict__ a, const Fq12* __restrict__ b, uint8_t* __restrict__ eq, size_t n) {
    size_t i = gtid();
    if (i >= n) return;
    const uint4* x = reinterpret_cast<const uint4*>(a + i);
    const uint4* y = reinterpret_cast<const uint4*>(b + i);
    uint32_t d = 0;
#pragma unroll
    for (int w = 0; w < 24; w++) {
        uint4 u = x[w], v = y[w];
        d |= (u.x ^ v.x) | (u.y ^ v.y) | (u.z ^ v.z) | (u.w ^ v.w);
    }
    eq[i] = d == 0;
}
template <class F>
BN_KERNEL k_jac_mul(const Jac<F>* __restrict__ p, const Fq* __restrict__ fr, Jac<F>* __restrict__ out, size_t n) {
    size_t i = gtid();
    if (i >= n) return;
    Jac<F> x = p[i], r;
    Fq m = fr[i];
    uint32_t k[8];
    fr_to_canonical(k, m.v);
    jac_mul(r, x, k);
    out[i] = r;
}
// Elementwise group operations (src/groups/mod.rs:250-269 double, :294-334 Add, :336-350 Neg, :169-195 PartialEq):
// op 0 = a + b, 1 = 2 a, 2 = -a, 3 = (a == b) as one byte per element
template <class F>
BN_KERNEL k_jac_elementwise(int op, const Jac<F>* __restrict__ a, const Jac<F>* __restrict__ b, Jac<F>* __restrict__ out,
                            uint8_t* __restrict__ eq, size_t n) {
    size_t i = gtid();
    if (i >= n) return;
    Jac<F> x = a[i], r;
    if (op == 0) { Jac<F> y = b[i]; jac_add(r, x, y); out[i] = r; }
    else if (op == 1) { jac_double(r, x); out[i] = r; }
    else if (op == 2) { jac_neg(r, x); out[i] = r; }
    else { Jac<F> y = b[i]; eq[i] = jac_eq(x, y) ? 1 : 0; }
}
// G * Fr on the reference's exact add/double sequence (src/groups/mod.rs:272-292; raw Jacobian bytes match), with the additions
// of a BLOCK compacted onto full warps.  k_jac_mul above gives every thread its own `if (bit) add`: with 32 independent scalars in
// a warp the addition runs for almost every bit with half the lanes idle (0.54 of the IMAD peak, DESIGN F16).  Here the block's
// running points and bases live in shared memory; per bit every thread doubles its own point, the threads whose bit is set
// append their index to a list (ballot + prefix sum), and the first `count` threads of the block perform the additions of the
// listed items -- about half the warps take the addition, fully occupied, the others skip it.  Each item still sees exactly the
// reference's sequence (double, then add when its bit is set), so the result is the same Jacobian triple.
#define BN_CMUL_BLOCK 256
template <class F> struct CmulLayout {
    static constexpr unsigned PT = sizeof(Jac<F>);              // 96 (G1) / 192 (G2)
    static constexpr unsigned STRIDE = 2 * PT + 16;              // res + base, odd number of 16-byte units: conflict-free per-thread access
    static constexpr unsigned LIST = BN_CMUL_BLOCK * STRIDE;     // uint32 list[BN_CMUL_BLOCK]
    static constexpr unsigned COUNTS = LIST + 4 * BN_CMUL_BLOCK; // uint32 warp_count[8] + total
    static constexpr unsigned BYTES = COUNTS + 64;
};
template <class F> __device__ __forceinline__ void cmul_load(Jac<F>& d, sp a) {
    uint4* w = reinterpret_cast<uint4*>(&d);
#pragma unroll
    for (unsigned k = 0; k < sizeof(Jac<F>) / 16; k++) w[k] = lds128_raw(a + 16 * k);
}
template <class F> __device__ __forceinline__ void cmul_store(sp a, const Jac<F>& v) {
    const uint4* w = reinterpret_cast<const uint4*>(&v);
#pragma unroll
    for (unsigned k = 0; k < sizeof(Jac<F>) / 16; k++) sts128_raw(a + 16 * k, w[k]);
}
template <class F>
__global__ void __launch_bounds__(BN_CMUL_BLOCK, 2) k_jac_mul_compact(const Jac<F>* __restrict__ p, const Fq* __restrict__ fr,
                                                                      Jac<F>* __restrict__ out, size_t n) {
    typedef CmulLayout<F> LY;
    const sp base = (sp)__cvta_generic_to_shared(bn_slot_mem);
    const unsigned tid = threadIdx.x, lane = tid & 31u, warp = tid >> 5;
    const size_t i = (size_t)blockIdx.x * BN_CMUL_BLOCK + tid;
    const bool live = i < n;
    uint32_t k[8];
    {
        Jac<F> x, z;
        jac_set_zero(z);
        if (live) {
            x = p[i];
            Fq m = fr[i];
            fr_to_canonical(k, m.v);
        } else {
            x = z;
#pragma unroll
            for (int j = 0; j < 8; j++) k[j] = 0;
        }
        cmul_store<F>(base + tid * LY::STRIDE, z);                  // running point: zero
        cmul_store<F>(base + tid * LY::STRIDE + LY::PT, x);         // base point
    }
    bool found_one = false;
    __syncthreads();
#pragma unroll 1
    for (int b = 255; b >= 0; b--) {
        const bool bit = (k[b >> 5] >> (b & 31)) & 1u;
        // every thread doubles its own running point (the reference doubles only after the first set bit)
        if (found_one) {
            Jac<F> r;
            cmul_load<F>(r, base + tid * LY::STRIDE);
            jac_double(r, r);
            cmul_store<F>(base + tid * LY::STRIDE, r);
        }
        found_one = found_one || bit;
        // list of the items whose bit is set
        const unsigned mask = __ballot_sync(0xffffffffu, bit);
        if (lane == 0) asm volatile("st.shared.u32 [%0], %1;" :: "r"(base + LY::COUNTS + 4u * warp), "r"(__popc(mask)) : "memory");
        __syncthreads();                                           // doublings stored, warp counts visible
        unsigned before = 0, total = 0;
#pragma unroll
        for (unsigned w = 0; w < BN_CMUL_BLOCK / 32; w++) {
            unsigned c;
            asm volatile("ld.shared.u32 %0, [%1];" : "=r"(c) : "r"(base + LY::COUNTS + 4u * w) : "memory");
            before += w < warp ? c : 0u;
            total += c;
        }
        if (bit) {
            const unsigned pos = before + __popc(mask & ((1u << lane) - 1u));
            asm volatile("st.shared.u32 [%0], %1;" :: "r"(base + LY::LIST + 4u * pos), "r"(tid) : "memory");
        }
        __syncthreads();                                           // list complete
        if (tid < total) {                                          // whole warps drop out: total is uniform
            unsigned item;
            asm volatile("ld.shared.u32 %0, [%1];" : "=r"(item) : "r"(base + LY::LIST + 4u * tid) : "memory");
            Jac<F> r, q;
            cmul_load<F>(r, base + item * LY::STRIDE);
            cmul_load<F>(q, base + item * LY::STRIDE + LY::PT);
            jac_add(r, r, q);
            cmul_store<F>(base + item * LY::STRIDE, r);
        }
        __syncthreads();                                           // additions stored before the next doubling; list / counts reusable
    }
    if (live) {
        Jac<F> r;
        cmul_load<F>(r, base + tid * LY::STRIDE);
        out[i] = r;
    }
}
// (p * k).normalize() by the signed 4-bit window (jac_mul_normalized, bn_curve.cuh); the 8-entry table lives in local memory
template <class F>
BN_KERNEL k_jac_mul_window(const Jac<F>* __restrict__ p, const Fq* __restrict__ fr, Jac<F>* __restrict__ out, size_t n) {
    size_t i = gtid();
    if (i >= n) return;
    Jac<F> x = p[i], r;
    Fq m = fr[i];
    uint32_t k[8];
    fr_to_canonical(k, m.v);
    jac_mul_normalized(r, x, k);
    out[i] = r;
}
template <class F>
BN_KERNEL k_jac_normalize(const Jac<F>* __restrict__ p, Jac<F>* __restrict__ out, size_t n) {
    size_t i = gtid();
    if (i >= n) return;
    Jac<F> x = p[i], r;
    jac_normalize(r, x);
    out[i] = r;
}

// AffineG::new (src/groups/mod.rs:94-113), SURVEY 8f N2: status 0 = Ok, 1 = NotOnCurve, 2 = NotInSubgroup.
// G1 has no order check (G1Params::check_order() == false, :401); G2 multiplies by r - 1 = -Fr::one() and adds p
// (:103), i.e. one full 254-bit G2 scalar multiplication per deserialised point.
BN_KERNEL k_g1_affine_check(const AffG1* __restrict__ pts, uint8_t* __restrict__ status, size_t n) {
    size_t i = gtid();
    if (i >= n) return;
    AffG1 a = pts[i];
    status[i] = g1_affine_status(a);
}
BN_KERNEL k_g2_affine_check(const AffG2* __restrict__ pts, uint8_t* __restrict__ status, size_t n, bool ref_ladder) {
    size_t i = gtid();
    if (i >= n) return;
    AffG2 a = pts[i];
    status[i] = g2_affine_status(a, ref_ladder);
}
// G1::from_compressed / G2::from_compressed (src/lib.rs:359-375, :506-526), SURVEY 8f N3.  33 / 65 bytes per
// record; a failed record writes an all-zero image and its CurveError as the status byte.
BN_KERNEL k_g1_from_compressed(const uint8_t* __restrict__ enc, G1* __restrict__ out, uint8_t* __restrict__ status, size_t n) {
    size_t i = gtid();
    if (i >= n) return;
    uint8_t rec[33];
    for (int j = 0; j < 33; j++) rec[j] = enc[i * 33 + j];
    G1 g;
    uint8_t st = g1_decode(g, rec);
    if (st != 0) { g.x = fq_zero(); g.y = fq_zero(); g.z = fq_zero(); }
    out[i] = g;
    status[i] = st;
}
BN_KERNEL k_g2_from_compressed(const uint8_t* __restrict__ enc, G2* __restrict__ out, uint8_t* __restrict__ status, size_t n,
                               bool ref_ladder) {
    size_t i = gtid();
    if (i >= n) return;
    uint8_t rec[65];
    for (int j = 0; j < 65; j++) rec[j] = enc[i * 65 + j];
    G2 g;
    uint8_t st = g2_decode(g, rec, ref_ladder);
    if (st != 0) { g.x = fq2_zero(); g.y = fq2_zero(); g.z = fq2_zero(); }
    out[i] = g;
    status[i] = st;
}

// field-level test hook (bn254_field_op_many)
BN_KERNEL k_field_op(int op, const uint4* __restrict__ a, const uint4* __restrict__ b, uint4* __restrict__ out, size_t n) {
    size_t i = gtid();
    if (i >= n) return;
    switch (op) {
        case 0: case 1: case 2: case 9: {
            Fq x = reinterpret_cast<const Fq*>(a)[i], r;
            if (op == 9) {   // all inversions must agree: Fermat power, its window, the uniform binary extended Euclid and its
                             // 30-step rounds, which must also end on their own (status 1, or 0 for zero) -- all-ones if not
                Fq r2, r3, r4;
                fq_inv(r, x); fq_inv_eea(r2, x); fq_inv_w4(r3, x);
                const int st = fq_inv_bgcd_core(r4, x);
                if (!fq_eq(r, r2) || !fq_eq(r, r3) || !fq_eq(r, r4) || st == 2) { for (int k = 0; k < 8; k++) r.v[k] = 0xffffffffu; }
            }
            else {
                Fq y = reinterpret_cast<const Fq*>(b)[i];
                if (op == 0) fq_mul(r, x, y); else if (op == 1) fq_add(r, x, y); else fq_sub(r, x, y);
            }
            reinterpret_cast<Fq*>(out)[i] = r;
        } break;
        case 3: case 4: case 10: {
            Fq2 x = reinterpret_cast<const Fq2*>(a)[i], r;
            if (op == 3) { Fq2 y = reinterpret_cast<const Fq2*>(b)[i]; fq2_mul(r, x, y); } else if (op == 4) fq2_sqr(r, x); else fq2_mul_by_xi(r, x);
            reinterpret_cast<Fq2*>(out)[i] = r;
        } break;
        case 5: {
            Fq6 x = reinterpret_cast<const Fq6*>(a)[i], y = reinterpret_cast<const Fq6*>(b)[i], r;
            fq6_mul(r, x, y);
            reinterpret_cast<Fq6*>(out)[i] = r;
        } break;
        default: {
            Fq12 x = reinterpret_cast<const Fq12*>(a)[i], r;
            if (op == 6) { Fq12 y = reinterpret_cast<const Fq12*>(b)[i]; fq12_mul(r, x, y); }
            else if (op == 7) fq12_sqr(r, x);
            else fq12_cyclotomic_sqr(r, x);
            reinterpret_cast<Fq12*>(out)[i] = r;
        } break;
    }
}

// slot-file test hook (ops 11..15 of bn254_field_op_many): the bodies of bn_slots.cuh one call at a time, launched with
// BN_BLOCK * BN_SLOT_STRIDE bytes of dynamic shared memory.  11: out[0] = sum_k a[k] * b[k] over three Fq2 pairs (s_dot3, unit
// 192 B, the other two output coefficients zero); 12: Fq12 product; 13: Fq12 squaring; 14: cyclotomic squaring; 15: Fq6 product.
BN_KERNEL k_slot_op(int op, const uint4* __restrict__ a, const uint4* __restrict__ b, uint4* __restrict__ out, size_t n) {
    size_t i = gtid();
    if (i >= n) return;
    const sp base = s_base();
    if (op == 11) {
        Fq12 z = fq12_zero();
        Fq6 x = reinterpret_cast<const Fq6*>(b)[i], r = fq6_zero();
        z.c0 = reinterpret_cast<const Fq6*>(a)[i];
        s_in(base, x.c0); s_in(base + 64, x.c1); s_in(base + 128, x.c2);
        s_dot3(r.c0, z, BN_PK3(0, 1, 2), base, BN_PK3(0, 1, 2));
        reinterpret_cast<Fq6*>(out)[i] = r;
    } else if (op == 15) {
        Fq6 x = reinterpret_cast<const Fq6*>(a)[i], y = reinterpret_cast<const Fq6*>(b)[i], r;
        s_fq6_mul(r, x, y, x, x, 0, base);
        reinterpret_cast<Fq6*>(out)[i] = r;
    } else {
        Fq12 x = reinterpret_cast<const Fq12*>(a)[i], r;
        if (op == 12) { Fq12 y = reinterpret_cast<const Fq12*>(b)[i]; s_fq12_mul_na(r, x, y, false, base); }
        else if (op == 13) s_fq12_sqr_na(r, x, base);
        else s_cyclotomic_sqr_na(r, x, base);
        reinterpret_cast<Fq12*>(out)[i] = r;
    }
}

// ---------------------------------------------------------------------------------------------
// roofline microbenchmarks: register-only chains.  kind 0: mad.lo.u32 (IMAD), 1: mad.wide.u32
// (IMAD.WIDE), 2: dependent Fq Montgomery multiplications (2 independent streams per thread).
// ---------------------------------------------------------------------------------------------
template <int KIND>
__global__ void __launch_bounds__(256) k_microbench(uint32_t* sink, int iters, uint32_t seed) {
    uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    if (KIND == 0) {
        uint32_t x[8];
#pragma unroll
        for (int j = 0; j < 8; j++) x[j] = t * 2654435761u + j + seed;
        uint32_t m = seed | 1u;
        for (int it = 0; it < iters; it++) {
#pragma unroll
            for (int r = 0; r < 16; r++) {
#pragma unroll
                for (int j = 0; j < 8; j++) asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(x[j]) : "r"(m), "r"(x[(j + 1) & 7]));
            }
        }
        uint32_t s = 0;
#pragma unroll
        for (int j = 0; j < 8; j++) s ^= x[j];
        if (s == 0x12345678u) sink[0] = s;
    } else if (KIND == 1) {
        unsigned long long x[8];
#pragma unroll
        for (int j = 0; j < 8; j++) x[j] = ((unsigned long long)t << 20) + j + seed;
        uint32_t m = seed | 1u;
        for (int it = 0; it < iters; it++) {
#pragma unroll
            for (int r = 0; r < 16; r++) {
#pragma unroll
                for (int j = 0; j < 8; j++) {
                    uint32_t lo = (uint32_t)x[j];
                    asm volatile("mad.wide.u32 %0, %1, %2, %0;" : "+l"(x[j]) : "r"(lo), "r"(m));
                }
            }
        }
        unsigned long long s = 0;
#pragma unroll
        for (int j = 0; j < 8; j++) s ^= x[j];
        if (s == 0x12345678ull) sink[0] = (uint32_t)s;
    } else {
        Fq a, b, c;
#pragma unroll
        for (int j = 0; j < 8; j++) { a.v[j] = t + j + seed; b.v[j] = t * 3 + j; c.v[j] = j + 1; }
        a.v[7] &= 0x0fffffffu; b.v[7] &= 0x0fffffffu;
        for (int it = 0; it < iters; it++) {
#pragma unroll
            for (int r = 0; r < 4; r++) { fq_mul(a, a, c); fq_mul(b, b, c); }
        }
        uint32_t s = 0;
#pragma unroll
        for (int j = 0; j < 8; j++) s ^= a.v[j] ^ b.v[j];
        if (s == 0x12345678u) sink[0] = s;
    }
}

}  // namespace bn
