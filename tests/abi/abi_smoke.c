/* A compiled C consumer of include/bn254_b200.h (no Python, no ctypes): what a Rust `extern "C"` block or a bindgen user
 * meets -- struct sizes, argument order, NULL-able outputs, status codes.  Runs the reference's own golden vectors
 * (test_reduced_pairing src/groups/mod.rs:928-999, test_miller_loop :642-691) and a 4-pair product through the header alone.
 * Built and run by tests/test_gpu_parity.py::test_abi_smoke_compiled_c_consumer. */
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "bn254_b200.h"
#include "golden_vectors.h"

#define CHECK(cond, what)                                                                   \
    do {                                                                                    \
        if (!(cond)) {                                                                      \
            fprintf(stderr, "abi_smoke FAILED: %s (%s)\n", what, bn254_last_error());       \
            return 1;                                                                       \
        }                                                                                   \
    } while (0)

int main(void) {
    /* the byte sizes the header promises are the crate's repr(C) sizes */
    CHECK(sizeof(VEC_G1_GEN) == BN254_G1_BYTES && sizeof(VEC_G2_GEN) == BN254_G2_BYTES, "G1/G2 sizes");
    CHECK(sizeof(VEC_GT_ONE) == BN254_GT_BYTES && sizeof(VEC_FR_G1_SCALAR) == BN254_FR_BYTES, "Gt/Fr sizes");
    CHECK(BN254_G2PRECOMP_BYTES == 128 + BN254_N_COEFFS * BN254_ELLCOEFFS_BYTES, "G2Precomp image size");

    bn254_ctx_options opt;
    bn254_ctx_options_default(&opt);
    CHECK(opt.struct_size == sizeof(bn254_ctx_options), "options struct size");
    bn254_ctx* ctx = NULL;
    CHECK(bn254_ctx_create_ex(0, &opt, &ctx) == BN254_OK && ctx != NULL, "ctx_create_ex");
    CHECK(bn254_ctx_device(ctx) == 0, "ctx_device");

    /* P = a G1, Q = b G2 with the reference's vector scalars (src/groups/mod.rs:646-653) */
    unsigned char p[BN254_G1_BYTES], q[BN254_G2_BYTES], gt[BN254_GT_BYTES], ml[BN254_GT_BYTES], fe[BN254_GT_BYTES];
    CHECK(bn254_g1_mul_many(ctx, VEC_G1_GEN, VEC_FR_G1_SCALAR, p, 1) == BN254_OK, "g1_mul_many");
    CHECK(bn254_g2_mul_many(ctx, VEC_G2_GEN, VEC_FR_G2_SCALAR, q, 1) == BN254_OK, "g2_mul_many");
    CHECK(bn254_pairing_many(ctx, p, q, gt, 1) == BN254_OK, "pairing_many");
    CHECK(memcmp(gt, VEC_GT_REDUCED_PAIRING, BN254_GT_BYTES) == 0, "test_reduced_pairing vector");
    /* miller_loop_batch takes (G2, G1) -- the reference's argument order, src/lib.rs:625 */
    CHECK(bn254_miller_loop_batch(ctx, q, p, 1, ml) == BN254_OK, "miller_loop_batch");
    CHECK(memcmp(ml, VEC_GT_MILLER_LOOP, BN254_GT_BYTES) == 0, "test_miller_loop vector");
    uint8_t some = 0;
    CHECK(bn254_final_exp_many(ctx, ml, fe, &some, 1) == BN254_OK && some == 1, "final_exp_many");
    CHECK(memcmp(fe, gt, BN254_GT_BYTES) == 0, "final_exponentiation(miller_loop) == pairing");

    /* a 4-pair product that is exactly one: e(P, Q) e(-P, Q) e(P, Q) e(P, -Q), with -P = (r-1) P and -Q = (r-1) Q */
    unsigned char pn[BN254_G1_BYTES], qn[BN254_G2_BYTES];
    CHECK(bn254_g1_mul_many(ctx, p, VEC_FR_MINUS_ONE, pn, 1) == BN254_OK, "g1_mul_many (-1)");
    CHECK(bn254_g2_mul_many(ctx, q, VEC_FR_MINUS_ONE, qn, 1) == BN254_OK, "g2_mul_many (-1)");
    unsigned char g1s[4 * BN254_G1_BYTES], g2s[4 * BN254_G2_BYTES], prod[BN254_GT_BYTES];
    memcpy(g1s, p, BN254_G1_BYTES); memcpy(g1s + BN254_G1_BYTES, pn, BN254_G1_BYTES);
    memcpy(g1s + 2 * BN254_G1_BYTES, p, BN254_G1_BYTES); memcpy(g1s + 3 * BN254_G1_BYTES, p, BN254_G1_BYTES);
    memcpy(g2s, q, BN254_G2_BYTES); memcpy(g2s + BN254_G2_BYTES, q, BN254_G2_BYTES);
    memcpy(g2s + 2 * BN254_G2_BYTES, q, BN254_G2_BYTES); memcpy(g2s + 3 * BN254_G2_BYTES, qn, BN254_G2_BYTES);
    uint8_t is_one = 0;
    CHECK(bn254_pairing_product_many(ctx, g1s, g2s, 4, 1, prod, &is_one) == BN254_OK, "pairing_product_many");
    CHECK(is_one == 1 && memcmp(prod, VEC_GT_ONE, BN254_GT_BYTES) == 0, "4-pair product == one");
    /* NULL-able outputs: only the accept byte / only the Gt */
    is_one = 7;
    CHECK(bn254_pairing_product_many(ctx, g1s, g2s, 4, 1, NULL, &is_one) == BN254_OK && is_one == 1, "gt_out = NULL");
    memset(prod, 0, sizeof prod);
    CHECK(bn254_pairing_product_many(ctx, g1s, g2s, 4, 1, prod, NULL) == BN254_OK, "is_one = NULL");
    CHECK(memcmp(prod, VEC_GT_ONE, BN254_GT_BYTES) == 0, "is_one = NULL result");
    /* three pairs only: not one */
    CHECK(bn254_pairing_product_many(ctx, g1s, g2s, 3, 1, prod, &is_one) == BN254_OK && is_one == 0, "3-pair product != one");
    CHECK(memcmp(prod, gt, BN254_GT_BYTES) == 0, "e(P,Q) e(-P,Q) e(P,Q) == e(P,Q)");
    uint8_t eq = 0;
    CHECK(bn254_gt_eq_many(ctx, prod, gt, &eq, 1) == BN254_OK && eq == 1, "gt_eq_many");

    /* the same product over a context list (two contexts on device 0) and the normalised scalar multiplication */
    bn254_ctx* ctx2 = NULL;
    CHECK(bn254_ctx_create(0, &ctx2) == BN254_OK, "second context");
    bn254_ctx* list[2] = {ctx, ctx2};
    is_one = 0;
    CHECK(bn254_pairing_product_sharded(list, 2, g1s, g2s, 4, prod, &is_one) == BN254_OK && is_one == 1, "pairing_product_sharded");
    unsigned char gts[4 * BN254_GT_BYTES], gts1[4 * BN254_GT_BYTES];
    CHECK(bn254_pairing_many_sharded(list, 2, g1s, g2s, gts, 4) == BN254_OK, "pairing_many_sharded");
    CHECK(bn254_pairing_many(ctx, g1s, g2s, gts1, 4) == BN254_OK && memcmp(gts, gts1, sizeof gts) == 0, "sharded == single");
    CHECK(memcmp(gts, gt, BN254_GT_BYTES) == 0, "sharded result 0 == golden");
    unsigned char pnorm[BN254_G1_BYTES], pnorm2[BN254_G1_BYTES];
    CHECK(bn254_g1_mul_normalized_many(ctx, VEC_G1_GEN, VEC_FR_G1_SCALAR, pnorm, 1) == BN254_OK, "g1_mul_normalized_many");
    CHECK(bn254_g1_normalize_many(ctx, p, pnorm2, 1) == BN254_OK && memcmp(pnorm, pnorm2, BN254_G1_BYTES) == 0, "normalised forms agree");

    /* error behaviour: infinity in miller_loop_batch -> BN254_ERR_TO_AFFINE, bad arguments -> negative codes */
    unsigned char inf[BN254_G1_BYTES];
    memset(inf, 0, sizeof inf);
    CHECK(bn254_miller_loop_batch(ctx, q, inf, 1, ml) == BN254_ERR_TO_AFFINE, "ToAffineConversion");
    CHECK(bn254_pairing_many(ctx, NULL, q, gt, 1) == BN254_ERR_ARG, "null pointer -> BN254_ERR_ARG");
    CHECK(bn254_pairing_many(NULL, p, q, gt, 1) == BN254_ERR_ARG, "null context -> BN254_ERR_ARG");
    CHECK(bn254_pairing_many(ctx, p, q, gt, 0) == BN254_OK, "n = 0 is a no-op");

    bn254_ctx_destroy(ctx2);
    bn254_ctx_destroy(ctx);
    printf("abi_smoke ok: %s\n", bn254_build_info());
    return 0;
}
