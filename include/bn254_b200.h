/*
 * bn254_b200 -- C ABI of the B200-native batched BN254 pairing engine.
 *
 * Drop-in boundary for the pairing surface of the `substrate-bn` crate (risc0/paritytech-bn).
 * The reference has no FFI of its own; what makes this boundary possible is that every public
 * type is #[repr(C)] down to U256([u128; 2]) (src/arith.rs:9-11), so the byte images below are
 * exactly the Rust values:
 *
 *   Fr, Fq   32 B   little-endian 256-bit MONTGOMERY residue (R = 2^256), src/lib.rs:16-18, :125-127
 *   Fq2      64 B   (c0, c1)                                              src/lib.rs:218-220
 *   G1       96 B   Jacobian (x, y, z) of Fq                              src/lib.rs:322-324
 *   G2      192 B   Jacobian (x, y, z) of Fq2                             src/lib.rs:469-471
 *   Gt      384 B   Fq12 = 2 x Fq6 = 2 x 3 x Fq2                          src/lib.rs:584-586
 *   G2Precomp image (export only): AffineG2 (128 B) + 87 x EllCoeffs{ell_0, ell_vw, ell_vv} (192 B)
 *            = 16,832 B, field order of src/groups/mod.rs:566-577
 *
 * Every entry point is batched (pointer + count); the scalar functions of src/lib.rs are the n = 1
 * calls.  Host-pointer entry points are synchronous: they stage through the context's grow-only device
 * arenas (no cudaMalloc on the steady-state path) and return when the results are in the caller's
 * buffers.  Host buffers may be pageable, page-locked, or page-locked in part (a cudaHostRegister of a sub-range: copies are
 * split at the boundary, and the zero-copy result path is taken only when the whole output extent is mapped).
 * `*_dev` entry points take DEVICE pointers (16-byte aligned, checked: BN254_ERR_ARG otherwise)
 * and enqueue on the context's stream.  They do not synchronise, with two documented exceptions:
 *   - the first call that needs a larger workspace than any call before it grows the arena
 *     (cudaStreamSynchronize + cudaFree + cudaMalloc, once);
 *   - bn254_pairing_product_*_dev with a NULL d_gt_out synchronise before returning (the temporary that
 *     receives the Gt values comes from the arena and must not be reused by the next call too early).
 * Limits of the product entry points: at most 8 shared G2Precomp tables per check; more than 8 variable
 * pairs per check cannot be combined with shared tables.
 *
 * Threading: a context is ONE device, two streams and three arenas.  Every entry point locks the
 * context for its duration, so concurrent callers of one context are safe but serialise; use one
 * context per host thread (or per GPU) for concurrency.  bn254_last_error() is thread-local.
 *
 * Return codes: 0 = ok; > 0 = reference-level condition (documented per function);
 *               < 0 = runtime failure (-1 bad argument, -2 CUDA error, -3 out of memory).
 * Nothing aborts or throws across this boundary (C++ exceptions are caught inside and mapped to
 * BN254_ERR_NOMEM / BN254_ERR_CUDA).  There is NO CPU fallback: without a CUDA device
 * bn254_ctx_create fails and nothing else can be called.
 */
#ifndef BN254_B200_H
#define BN254_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define BN254_FR_BYTES 32
#define BN254_FQ_BYTES 32
#define BN254_G1_BYTES 96
#define BN254_G2_BYTES 192
#define BN254_GT_BYTES 384
#define BN254_ELLCOEFFS_BYTES 192
#define BN254_N_COEFFS 87
#define BN254_G2PRECOMP_BYTES (128 + 87 * 192)

#define BN254_OK 0
#define BN254_ERR_TO_AFFINE 1 /* CurveError::ToAffineConversion, src/lib.rs:109-115, :629-630 */
#define BN254_ERR_ARG (-1)
#define BN254_ERR_CUDA (-2)
#define BN254_ERR_NOMEM (-3)

typedef struct bn254_ctx bn254_ctx;             /* one device + one stream + scratch */
typedef struct bn254_g2precomp bn254_g2precomp; /* device-resident G2Precomp tables (engine-side handle;
                                                   the reference's G2Precomp is private, src/lib.rs:7) */

/* ---- context ------------------------------------------------------------------------------- */
/* Runtime switches of a context as plain data (no hidden global state).  -1 in any field = the
 * library default; bn254_ctx_options_default() fills the struct with the defaults, letting the BN254_*
 * environment variables named below override them (A/B runs without recompiling).  bn254_ctx_create(d, out)
 * is bn254_ctx_create_ex(d, NULL, out), i.e. exactly those defaults. */
typedef struct bn254_ctx_options {
    uint32_t struct_size;  /* sizeof(bn254_ctx_options) of the caller's header (forward compatibility) */
    int32_t pipeline;      /* BN254_PIPELINE      host-buffer pairing path: 0 = one batch; 1 or 2 = two part batches on two streams
                              (copies overlap kernels); 3 .. 8 (default 4) = that many parts; beyond one wave of 448-thread
                              blocks the count is capped so that every part is itself a full wave */
    int32_t big_blocks;    /* BN254_BIG_BLOCKS    1: 512-thread blocks when they save waves; 2: for every batch > 512 (tests) */
    int32_t fused_miller;  /* BN254_FUSED_MILLER  0: affine -> line table -> Miller from table; 1: one fused kernel (A/B) */
    int32_t split_fexp;    /* BN254_SPLIT_FEXP    with fexp_smem = 0: 1 = easy + hard kernels, 0 = one kernel */
    int32_t fexp_smem;     /* BN254_FEXP_SMEM     2: hard part on the slot file, one phase-synchronised block per SM; 1: 64-thread blocks; 0: local memory */
    int32_t miller_smem;   /* BN254_MILLER_SMEM   1: Miller loop from the table on the slot file; 0: local memory */
    int32_t lines_smem;    /* BN254_LINES_SMEM    1: G2 line tables on the slot file (one-wave batches); 0: local memory */
    int32_t subgroup_ref;  /* BN254_SUBGROUP_REF  0: G2 order check by the psi test; 1: the reference's [r-1]P + P ladder.
                              Both decide the same predicate; this picks the algorithm that validates untrusted input. */
    int32_t zerocopy_out;  /* BN254_ZEROCOPY_OUT  1: pinned (mapped) output buffers are written by the last kernel directly */
    int32_t coop_max_n;    /* BN254_COOP_MAX_N    batches of at most this many pairings take the warp-cooperative latency path
                              (one warp per pairing); 0 disables it */
    int32_t stack_bytes;   /* BN254_STACK_BYTES   0: leave cudaLimitStackSize alone (the driver sizes static call graphs itself);
                              > 0: raise the DEVICE-WIDE limit to at least this (affects the embedding process) */
    int32_t mul_compact;   /* BN254_MUL_COMPACT   1: G1 * Fr (reference sequence) compacts a block's additions onto full warps
                              (k_jac_mul_compact, +6 %); 2: G2 * Fr as well (no gain, tests); 0: one thread per scalar with its
                              own `if (bit) add` (k_jac_mul).  Same Jacobian bytes in every mode. */
    int32_t coop4_max_n;   /* BN254_COOP4_MAX_N   batches of at most this many units (default 592 = 4 per SM) take the quad-lane
                              executor of the cooperative microcode: four warps per pairing, lowest latency; 0 disables */
} bn254_ctx_options;
void bn254_ctx_options_default(bn254_ctx_options* opt);
int bn254_ctx_create(int device, bn254_ctx** out);
int bn254_ctx_create_ex(int device, const bn254_ctx_options* opt, bn254_ctx** out);
void bn254_ctx_destroy(bn254_ctx* ctx);
int bn254_ctx_device(const bn254_ctx* ctx);
/* the CUDA stream (cudaStream_t) all work of this context is enqueued on */
void* bn254_ctx_stream(const bn254_ctx* ctx);
int bn254_ctx_synchronize(bn254_ctx* ctx);
/* number of kernels this context has launched so far (bench.py's gpu_launches evidence) */
uint64_t bn254_ctx_launch_count(const bn254_ctx* ctx);
const char* bn254_last_error(void);
const char* bn254_build_info(void);

/* ---- pairing: src/lib.rs:611-613 `pairing(p: G1, q: G2) -> Gt`, n independent pairs ------------ */
/* A pair with either point at infinity yields Gt::one() (src/groups/mod.rs:896). */
int bn254_pairing_many(bn254_ctx* ctx, const void* g1, const void* g2, void* gt_out, size_t n);
int bn254_pairing_many_dev(bn254_ctx* ctx, const void* d_g1, const void* d_g2, void* d_gt_out, size_t n);
/* The same over a device list (SURVEY 8b "device-list argument", 8e): ONE call, n independent pairs in one host
 * buffer, contiguous index ranges [n*d/n_ctx, n*(d+1)/n_ctx) on contexts ctxs[0..n_ctx) (one per GPU; several
 * contexts on one GPU are allowed), one host thread and stream pair per device, every result written to the caller's
 * own index range of gt_out.  No data-path collective: the pairs are independent.  Pinned (cudaHostAlloc /
 * cudaHostRegister) buffers make the copies asynchronous; pageable buffers work and are staged by the driver. */
int bn254_pairing_many_sharded(bn254_ctx** ctxs, size_t n_ctx, const void* g1, const void* g2, void* gt_out, size_t n);

/* Profiling twin of bn254_pairing_many_dev: identical launches, each kernel bracketed by CUDA events on the
 * context's stream.  ms[5] = {affine conversion, G2 line table, Miller loop from table, easy part of the final
 * exponentiation, hard part} in milliseconds (fused-Miller option: ms[0] = ms[1] = 0, ms[2] = the fused kernel;
 * single-kernel final exponentiation: ms[3] = 0, ms[4] = that kernel).  Synchronises. */
int bn254_pairing_many_dev_timed(bn254_ctx* ctx, const void* d_g1, const void* d_g2, void* d_gt_out, size_t n, float* ms);

/* ---- pairing_batch: src/lib.rs:615-623 `pairing_batch(&[(G1, G2)]) -> Gt` ---------------------- */
/* n_checks independent products of pairs_per_check pairs each (pair j of check i at index
 * i*pairs_per_check + j), ONE final exponentiation per check.  Pairs with a point at infinity are
 * skipped; a check with no remaining pair yields one (src/groups/mod.rs:904-926).
 * gt_out (n_checks x 384 B) and is_one (n_checks bytes: result == Gt::one()) may each be NULL. */
int bn254_pairing_product_many(bn254_ctx* ctx, const void* g1, const void* g2, size_t pairs_per_check,
                               size_t n_checks, void* gt_out, uint8_t* is_one);
int bn254_pairing_product_many_dev(bn254_ctx* ctx, const void* d_g1, const void* d_g2, size_t pairs_per_check,
                                   size_t n_checks, void* d_gt_out, uint8_t* d_is_one);

/* ---- miller_loop_batch: src/lib.rs:625-633 `miller_loop_batch(&[(G2, G1)]) -> Result<Gt, _>` ---- */
/* NOTE the (G2, G1) order.  One un-exponentiated product over all n pairs.  Returns
 * BN254_ERR_TO_AFFINE if any point is at infinity (gt_out untouched), like the reference. */
int bn254_miller_loop_batch(bn254_ctx* ctx, const void* g2, const void* g1, size_t n, void* gt_out);
/* n independent single-pair Miller loops (pre-final-exponentiation values); ok[i] = 0 and gt = one
 * where a point is at infinity.  ok may be NULL. */
int bn254_miller_loop_many(bn254_ctx* ctx, const void* g2, const void* g1, void* gt_out, uint8_t* ok, size_t n);
int bn254_miller_loop_many_dev(bn254_ctx* ctx, const void* d_g2, const void* d_g1, void* d_gt_out, uint8_t* d_ok,
                               size_t n);

/* Un-exponentiated product of the Miller values of n pairs with pairing_batch's skip semantics (a pair with a
 * point at infinity contributes one; n = 0 yields one).  This is the per-GPU partial of a product sharded over
 * several processes / devices: partials multiply exactly (src/groups/mod.rs:609-640). */
int bn254_miller_product(bn254_ctx* ctx, const void* g1, const void* g2, size_t n, void* gt_out);

/* ---- Gt: src/lib.rs:584-609 -------------------------------------------------------------------- */
/* Gt::final_exponentiation -> Option<Gt>: some[i] = 0 (and gt_out = 0) iff the input is zero */
int bn254_final_exp_many(bn254_ctx* ctx, const void* gt_in, void* gt_out, uint8_t* some, size_t n);
int bn254_final_exp_many_dev(bn254_ctx* ctx, const void* d_gt_in, void* d_gt_out, uint8_t* d_some, size_t n);
/* Same as bn254_final_exp_many_dev for n <= BN254_NOALLOC_MAX_N, allocating NOTHING: no cudaMalloc, cudaFree or allocating
 * library routine on any call.  No counterpart in the reference: added for callers
 * that cannot call cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it).  The easy part's
 * intermediate values live in a static device array of the library (BN254_NOALLOC_MAX_N x 385 bytes = 1.5 MB, held by every
 * GPU that loads the library until exit) instead of the context's grow-only workspace.  n > BN254_NOALLOC_MAX_N returns
 * BN254_ERR_ARG.  All calls on one device share that array: calls that overlap on the GPU (two contexts of one device) race;
 * calls of one context are stream-ordered and safe.  Enqueues on the context's stream without synchronising; d_gt_out may
 * alias d_gt_in; d_some may be NULL. */
#define BN254_NOALLOC_MAX_N 4096
int bn254_final_exp_many_dev_noalloc(bn254_ctx* ctx, const void* d_gt_in, void* d_gt_out, uint8_t* d_some, size_t n);
int bn254_gt_mul_many(bn254_ctx* ctx, const void* a, const void* b, void* out, size_t n);     /* Gt * Gt   */
int bn254_gt_pow_many(bn254_ctx* ctx, const void* a, const void* fr, void* out, size_t n);    /* Gt::pow   */
int bn254_gt_inverse_many(bn254_ctx* ctx, const void* a, void* out, uint8_t* some, size_t n); /* Gt::inverse */
int bn254_gt_eq_many(bn254_ctx* ctx, const void* a, const void* b, uint8_t* eq, size_t n);    /* PartialEq */

/* ---- G1 * Fr, G2 * Fr: src/lib.rs:425-431, :576-582 -------------------------------------------- */
/* Fr is the crate's Montgomery image; the kernel de-Montgomeries it (src/groups/mod.rs:279).
 * The result is the reference's own Jacobian representative (same add/double sequence). */
int bn254_g1_mul_many(bn254_ctx* ctx, const void* g1, const void* fr, void* g1_out, size_t n);
int bn254_g2_mul_many(bn254_ctx* ctx, const void* g2, const void* fr, void* g2_out, size_t n);
int bn254_g1_mul_many_dev(bn254_ctx* ctx, const void* d_g1, const void* d_fr, void* d_out, size_t n);
int bn254_g2_mul_many_dev(bn254_ctx* ctx, const void* d_g2, const void* d_fr, void* d_out, size_t n);
/* (p * k).normalize() in one call for callers that normalise anyway (src/lib.rs:425-431 followed by :391-398): a signed
 * 4-bit window with uniform control flow (64 table additions instead of the ~254 a warp executes on the bit-serial
 * ladder).  Normalised coordinates are unique, so the bytes equal the reference's; a result at infinity (k = 0, p at
 * infinity, or a non-member whose multiple vanishes) is recomputed with the reference's own ladder so that even those
 * bytes match. */
int bn254_g1_mul_normalized_many(bn254_ctx* ctx, const void* g1, const void* fr, void* g1_out, size_t n);
int bn254_g2_mul_normalized_many(bn254_ctx* ctx, const void* g2, const void* fr, void* g2_out, size_t n);
int bn254_g1_mul_normalized_many_dev(bn254_ctx* ctx, const void* d_g1, const void* d_fr, void* d_out, size_t n);
int bn254_g2_mul_normalized_many_dev(bn254_ctx* ctx, const void* d_g2, const void* d_fr, void* d_out, size_t n);
/* Group::normalize, src/lib.rs:391-398, :542-549: z -> one (infinity untouched) */
int bn254_g1_normalize_many(bn254_ctx* ctx, const void* g1, void* g1_out, size_t n);
int bn254_g2_normalize_many(bn254_ctx* ctx, const void* g2, void* g2_out, size_t n);

/* ---- the group law itself, elementwise: src/lib.rs:405-423, :556-574 (Add, Sub via Neg, Neg), Group::double, derived ==
 * -> src/groups/mod.rs:294-334 (add-2007-bl with the zero / equal-point branches), :250-269 (dbl-2009-l), :336-350 (Neg),
 * :169-195 (PartialEq: PROJECTIVE equality, eq[i] = 1 when both triples name the same point).  Results are the reference's own
 * Jacobian representatives (same formulas), so raw bytes match. */
int bn254_g1_add_many(bn254_ctx* ctx, const void* a, const void* b, void* out, size_t n);
int bn254_g2_add_many(bn254_ctx* ctx, const void* a, const void* b, void* out, size_t n);
int bn254_g1_double_many(bn254_ctx* ctx, const void* a, void* out, size_t n);
int bn254_g2_double_many(bn254_ctx* ctx, const void* a, void* out, size_t n);
int bn254_g1_neg_many(bn254_ctx* ctx, const void* a, void* out, size_t n);
int bn254_g2_neg_many(bn254_ctx* ctx, const void* a, void* out, size_t n);
int bn254_g1_eq_many(bn254_ctx* ctx, const void* a, const void* b, uint8_t* eq, size_t n);
int bn254_g2_eq_many(bn254_ctx* ctx, const void* a, const void* b, uint8_t* eq, size_t n);

/* ---- AffineG1::new / AffineG2::new: src/lib.rs:437-447, :639-649 -> src/groups/mod.rs:94-113 ---- */
/* xy: n affine points (x, y) as Montgomery images (64 B for G1, 128 B for G2).  status[i] = 0 Ok,
 * 1 GroupError::NotOnCurve, 2 GroupError::NotInSubgroup (G2 only: p * (-1) + p != 0). */
#define BN254_AFFINE_OK 0
#define BN254_AFFINE_NOT_ON_CURVE 1
#define BN254_AFFINE_NOT_IN_SUBGROUP 2
int bn254_g1_affine_check_many(bn254_ctx* ctx, const void* xy, uint8_t* status, size_t n);
int bn254_g2_affine_check_many(bn254_ctx* ctx, const void* xy, uint8_t* status, size_t n);
/* device-pointer twins (asynchronous on the context's stream): validate points that are already in HBM */
int bn254_g1_affine_check_many_dev(bn254_ctx* ctx, const void* d_xy, uint8_t* d_status, size_t n);
int bn254_g2_affine_check_many_dev(bn254_ctx* ctx, const void* d_xy, uint8_t* d_status, size_t n);

/* ---- G1::from_compressed src/lib.rs:359-375, G2::from_compressed :506-526 ---- */
/* enc: n records of 33 bytes (G1: tag 2/3, x big-endian) or 65 bytes (G2: tag 10/11, x = c1*q + c0 as a
 * 512-bit big-endian integer).  out: n Jacobian images with z = one (all-zero bytes for a failed record).
 * status[i] = the reference's Result: 0 Ok, 1 CurveError::InvalidEncoding (tag), 2 CurveError::NotMember
 * (no square root, or AffineG::new refused the point), 3 CurveError::Field(FieldError::NotMember) (x >= q).
 * The checks run in the reference's order: field, square root, tag, AffineG::new.  The length check of the
 * slice (:360, :508) is the caller's: the record size is fixed by this ABI. */
#define BN254_DECODE_OK 0
#define BN254_DECODE_INVALID_ENCODING 1
#define BN254_DECODE_NOT_MEMBER 2
#define BN254_DECODE_FIELD_NOT_MEMBER 3
int bn254_g1_from_compressed_many(bn254_ctx* ctx, const uint8_t* enc, void* g1_out, uint8_t* status, size_t n);
int bn254_g2_from_compressed_many(bn254_ctx* ctx, const uint8_t* enc, void* g2_out, uint8_t* status, size_t n);
/* device-pointer twins: the decoded images stay in HBM and can be handed to bn254_pairing_many_dev and friends
 * (a failed record is all-zero, i.e. z = 0: the pairing entry points treat it as the point at infinity) */
int bn254_g1_from_compressed_many_dev(bn254_ctx* ctx, const uint8_t* d_enc, void* d_g1_out, uint8_t* d_status, size_t n);
int bn254_g2_from_compressed_many_dev(bn254_ctx* ctx, const uint8_t* d_enc, void* d_g2_out, uint8_t* d_status, size_t n);

/* ---- shared G2Precomp: src/groups/mod.rs:573-577, :701-728 ------------------------------------- */
/* Precompute the 87 line-coefficient triples of n G2 points once; the tables stay on the device.
 * Returns BN254_ERR_TO_AFFINE if a point is at infinity. */
int bn254_g2_precompute(bn254_ctx* ctx, const void* g2, size_t n, bn254_g2precomp** out);
void bn254_g2precomp_destroy(bn254_g2precomp* pc);
size_t bn254_g2precomp_count(const bn254_g2precomp* pc);
/* copy table i back in the reference's G2Precomp field order (BN254_G2PRECOMP_BYTES) */
int bn254_g2precomp_export(bn254_ctx* ctx, const bn254_g2precomp* pc, size_t i, void* out);
/* Groth16-shaped product checks with shared tables: check i multiplies
 *   prod_{j < m}  e(g1_shared[i*m + j], table j)   (m = bn254_g2precomp_count(pc))
 *   prod_{j < v}  e(g1_var[i*v + j], g2_var[i*v + j])
 * with one shared squaring chain and one final exponentiation.  Infinity pairs are skipped. */
int bn254_pairing_product_precomp_many(bn254_ctx* ctx, const bn254_g2precomp* pc, const void* g1_shared,
                                       const void* g1_var, const void* g2_var, size_t v, size_t n_checks,
                                       void* gt_out, uint8_t* is_one);
int bn254_pairing_product_precomp_many_dev(bn254_ctx* ctx, const bn254_g2precomp* pc, const void* d_g1_shared,
                                           const void* d_g1_var, const void* d_g2_var, size_t v, size_t n_checks,
                                           void* d_gt_out, uint8_t* d_is_one);

/* ---- one product sharded over several devices -------------------------------------- */
/* prod_i e(g1[i], g2[i]) over n pairs split by index range across n_ctx contexts (one per GPU):
 * each device reduces its range to one un-exponentiated Fq12 (384 B), the partials are gathered to
 * ctxs[0] by peer copy, multiplied, and exponentiated once.  Same skip semantics as pairing_batch. */
int bn254_pairing_product_sharded(bn254_ctx** ctxs, size_t n_ctx, const void* g1, const void* g2, size_t n,
                                  void* gt_out, uint8_t* is_one);
/* Same with the phases timed on the host clock (extra synchronisations, for bench.py): phase_ms[0] = copies in + Miller
 * partial products on all devices (wall, i.e. max over devices), [1] = gather of the n_ctx 384-byte partials to ctxs[0]
 * (peer copies), [2] = product of the partials + the one final exponentiation + result copy. */
int bn254_pairing_product_sharded_timed(bn254_ctx** ctxs, size_t n_ctx, const void* g1, const void* g2, size_t n,
                                        void* gt_out, uint8_t* is_one, double* phase_ms);
/* Device-resident twin of bn254_miller_product: d_gt_out receives the un-exponentiated product of n pairs already in
 * HBM (asynchronous on the context's stream).  The per-rank step of a product sharded over processes. */
int bn254_miller_product_dev(bn254_ctx* ctx, const void* d_g1, const void* d_g2, size_t n, void* d_gt_out);

/* ---- field-level test hooks (parity tests of the arithmetic kernels) ---------------------------- */
/* op: 0 = Fq mul, 1 = Fq add, 2 = Fq sub, 3 = Fq2 mul, 4 = Fq2 sqr, 5 = Fq6 mul, 6 = Fq12 mul,
 *     7 = Fq12 sqr, 8 = Fq12 cyclotomic_squared, 9 = Fq inverse (b ignored),
 *     10 = Fq2 mul_by_nonresidue (xi = 9 + u, src/fields/fq2.rs:55-57; b ignored);
 *     the shared-memory slot-file bodies one call at a time: 11 = sum of three Fq2 products a[k] * b[k] (unit 192 B, result in
 *     the first 64 B), 12 = Fq12 mul, 13 = Fq12 sqr, 14 = cyclotomic_squared, 15 = Fq6 mul */
int bn254_field_op_many(bn254_ctx* ctx, int op, const void* a, const void* b, void* out, size_t n);

/* ---- roofline microbenchmarks (register-only; report via bench.py) ------------------------------ */
/* kind: 0 = IMAD (mad.lo.u32) chains, 1 = IMAD.WIDE (mad.wide.u32) chains, 2 = back-to-back Fq
 * Montgomery multiplications.  Returns elapsed milliseconds for `iters` iterations x `ilp`
 * independent chains per thread over grid*block threads; *ops receives the op count. */
int bn254_microbench(bn254_ctx* ctx, int kind, int grid, int block, int iters, double* ms, double* ops);

#ifdef __cplusplus
}
#endif
#endif /* BN254_B200_H */
