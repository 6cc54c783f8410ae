"""Parity tests proper: the CUDA path, called through the C ABI (libbn254_b200.so via ctypes), against
the C oracle on the same seeded inputs, against the reference's golden vectors, and -- at the full
BASELINE.json sizes -- through size-independent properties.  Bit-exact everywhere (integer work)."""
import os

import numpy as np
import pytest

from oracle import bn254_ref as o
from oracle import c_oracle as c
from tests import util

pytestmark = pytest.mark.gpu

G1, G2, GT, FR = 96, 192, 384, 32


@pytest.fixture(scope="module")
def eng():
    import paritytech_bn_b200 as bn
    e = bn.Engine(0)
    yield e
    e.close()


def _b(x):
    return np.frombuffer(o.to_bytes(x), dtype=np.uint8).copy()


def _rand_fq_elems(rng, n_fq, n, edge=True):
    vals = rng.integers(0, 2**63, size=(n * n_fq, 5), dtype=np.uint64)
    out = []
    for row in vals:
        v = 0
        for w in row:
            v = (v << 63) | int(w)
        out.append(v % o.Q)
    if edge:
        for k, v in enumerate([0, 1, o.Q - 1, o.Q - 2, 2, o.MONT_R % o.Q]):
            if k < len(out):
                out[k * max(1, len(out) // 7)] = v
    return np.frombuffer(b"".join(o.fq_to_mont(v).to_bytes(32, "little") for v in out), dtype=np.uint8).copy()


@pytest.mark.parametrize("name,unit,ref", [
    ("fq_mul", 32, c.fq_mul), ("fq_add", 32, c.fq_add), ("fq_sub", 32, c.fq_sub),
    ("fq2_mul", 64, c.fq2_mul), ("fq6_mul", 192, c.fq6_mul), ("fq12_mul", 384, c.fq12_mul),
])
def test_field_binary(eng, name, unit, ref):
    rng = np.random.default_rng(1)
    n = 4096 if unit <= 64 else 512
    a, b = _rand_fq_elems(rng, unit // 32, n), _rand_fq_elems(rng, unit // 32, n)
    assert eng.field_op(name, a, b).tobytes() == ref(a, b).tobytes()
    assert eng.field_op(name, a, a).tobytes() == ref(a, a).tobytes()


@pytest.mark.parametrize("name,unit,ref", [
    ("fq2_sqr", 64, c.fq2_sqr), ("fq12_sqr", 384, c.fq12_sqr),
])
def test_field_unary(eng, name, unit, ref):
    rng = np.random.default_rng(2)
    a = _rand_fq_elems(rng, unit // 32, 512)
    assert eng.field_op(name, a).tobytes() == ref(a).tobytes()


def test_fq_inverse(eng):
    rng = np.random.default_rng(3)
    a = _rand_fq_elems(rng, 1, 256)
    ro, some = c.fq_inv(a)
    assert eng.field_op("fq_inv", a).tobytes() == ro.tobytes()     # zero -> zero


def test_golden_reduced_pairing(eng, golden):
    """src/groups/mod.rs:928-999 through the C ABI, incl. G*Fr and the pre-final-exp value (:642-691)."""
    v = golden["test_reduced_pairing"]
    k1 = util.fr_bytes([int(v["g1_scalar"])])
    k2 = util.fr_bytes([int(v["g2_scalar"])])
    p = eng.g1_mul(util.G1_GEN, k1)
    q = eng.g2_mul(util.G2_GEN, k2)
    assert p.tobytes() == c.g1_mul(util.G1_GEN, k1).tobytes()
    assert q.tobytes() == c.g2_mul(util.G2_GEN, k2).tobytes()
    gt = eng.pairing(p, q)
    assert gt.tobytes() == util.fq_mont_bytes([int(x) for x in v["expected"]])
    m = golden["test_miller_loop"]
    ml = eng.miller_loop_batch(q, p)
    assert ml.tobytes() == util.fq_mont_bytes([int(x) for x in m["expected"]])
    fe, some = eng.final_exponentiation(ml)
    assert some[0] == 1 and fe.tobytes() == gt.tobytes()


def test_golden_prepared_g2(eng, golden):
    """All 87 EllCoeffs of src/groups/mod.rs:779-892 from the device table."""
    v = golden["test_prepared_g2"]
    q = eng.g2_mul(util.G2_GEN, util.fr_bytes([int(v["g2_scalar"])]))
    pc = eng.g2_precompute(np.concatenate([q, util.G2_GEN]))
    assert len(pc) == 2
    exp = util.fq_mont_bytes([int(x) for x in v["q"]["x"] + v["q"]["y"]])
    for e in v["coeffs"]:
        exp += util.fq_mont_bytes([int(x) for x in e["ell_0"] + e["ell_vw"] + e["ell_vv"]])
    assert pc.export(0).tobytes() == exp
    assert pc.export(1).tobytes() == c.g2_precompute(util.G2_GEN).tobytes()
    pc.close()


def _workload(n, stream):
    ks1 = util.scalars(n, stream=stream)
    ks2 = util.scalars(n, stream=stream + 100)
    return util.points_cpu(ks1, ks2)


def test_pairing_many_vs_oracle(eng):
    n = int(os.environ.get("BN_TEST_PAIRINGS", "2048"))
    g1, g2 = _workload(n, 1)
    g1 = g1.reshape(n, G1).copy()
    g2 = g2.reshape(n, G2).copy()
    # injected edge cases: infinity on either side, generators with z == 1, a = r-1
    g1[3], g2[5] = util.G1_INF, util.G2_INF
    g1[7], g2[7] = util.G1_INF, util.G2_INF
    g1[9], g2[9] = util.G1_GEN, util.G2_GEN
    g1[11] = c.g1_mul(util.G1_GEN, util.fr_bytes([o.R_ORDER - 1]))
    g1, g2 = g1.reshape(-1), g2.reshape(-1)
    got = eng.pairing(g1, g2)
    exp = c.pairing_many(g1, g2)
    assert got.tobytes() == exp.tobytes()
    assert got.reshape(n, GT)[3].tobytes() == util.GT_ONE.tobytes()
    # canonical (non-Montgomery big-endian) serialisation agrees too
    can = c.fq_to_canonical(got[:GT])
    assert can.tobytes() == c.fq_to_canonical(exp[:GT]).tobytes()
    # the split public API: miller_loop_many + final_exponentiation
    ml, ok = eng.miller_loop_many(g2, g1)
    rml, rok = c.miller_loop_many(g2, g1)
    assert (ok == rok).all() and ml.tobytes() == rml.tobytes()
    fe, some = eng.final_exponentiation(ml)
    assert some.all() and fe.tobytes() == exp.tobytes()


def test_pairing_batch_checks_vs_oracle(eng):
    n_checks = 256
    w = util.groth16_checks(n_checks)
    out, is_one = eng.pairing_batch(w["g1"], w["g2"], pairs_per_check=4)
    rout, ris = c.pairing_product_many(w["g1"], w["g2"], 4)
    assert (is_one == w["valid"]).all() and (ris == is_one).all()
    assert out.tobytes() == rout.tobytes()
    # same checks through the shared-G2Precomp entry point
    pc = eng.g2_precompute(w["vk_g2"])
    out2, is_one2 = eng.pairing_batch_precomp(pc, w["g1_shared"], w["g1_var"], w["g2_var"], 1)
    assert out2.tobytes() == rout.tobytes() and (is_one2 == is_one).all()
    pc.close()


def test_pairing_batch_edge_semantics(eng):
    g1, g2 = _workload(6, 3)
    g1 = g1.reshape(6, G1).copy()
    g2 = g2.reshape(6, G2).copy()
    g1[1] = util.G1_INF          # skipped pair
    g2[2] = util.G2_INF          # skipped pair
    g1, g2 = g1.reshape(-1), g2.reshape(-1)
    out, _ = eng.pairing_batch(g1, g2)                      # one product of 6 pairs
    assert out.tobytes() == c.pairing_batch(g1, g2).tobytes()
    out3, _ = eng.pairing_batch(g1, g2, pairs_per_check=3)  # two checks of 3
    assert out3.tobytes() == c.pairing_product_many(g1, g2, 3)[0].tobytes()
    out2, _ = eng.pairing_batch(g1, g2, pairs_per_check=2)
    assert out2.tobytes() == c.pairing_product_many(g1, g2, 2)[0].tobytes()
    # all pairs skipped -> one
    out, is_one = eng.pairing_batch(np.tile(util.G1_INF, 2), g2[:2 * G2])
    assert out.tobytes() == util.GT_ONE.tobytes() and is_one[0] == 1
    # more pairs per check than the fused kernel handles (falls to per-pair Miller + product)
    g1b, g2b = _workload(24, 4)
    outb, _ = eng.pairing_batch(g1b, g2b, pairs_per_check=12)
    assert outb.tobytes() == c.pairing_product_many(g1b, g2b, 12)[0].tobytes()


def test_miller_loop_batch_api(eng):
    import paritytech_bn_b200 as bn
    g1, g2 = _workload(5, 5)
    assert eng.miller_loop_batch(g2, g1).tobytes() == c.miller_loop_batch(g2, g1).tobytes()
    assert eng.miller_loop_batch(g2[:G2], g1[:G1]).tobytes() == c.miller_loop_batch(g2[:G2], g1[:G1]).tobytes()
    assert eng.miller_loop_batch(np.zeros(0, np.uint8), np.zeros(0, np.uint8)).tobytes() == util.GT_ONE.tobytes()
    bad = g2.copy()
    bad[G2:2 * G2] = util.G2_INF
    with pytest.raises(bn.CurveError):
        eng.miller_loop_batch(bad, g1)


def test_scalar_mul_vs_oracle(eng):
    n = 1024
    ks = util.scalars(n, stream=11)
    base1, base2 = _workload(n, 12)
    sc = [0, 1, 2, o.R_ORDER - 1] + ks[4:]
    fr = util.fr_bytes(sc)
    got1 = eng.g1_mul(base1, fr)
    assert got1.tobytes() == c.g1_mul(base1, fr).tobytes()          # raw Jacobian equality
    n2 = 256
    got2 = eng.g2_mul(base2[:n2 * G2], fr[:n2 * FR])
    assert got2.tobytes() == c.g2_mul(base2[:n2 * G2], fr[:n2 * FR]).tobytes()
    assert eng.g1_normalize(got1).tobytes() == c.g1_normalize(got1).tobytes()
    assert eng.g2_normalize(got2).tobytes() == c.g2_normalize(got2).tobytes()
    inf = np.concatenate([util.G1_INF, util.G1_GEN])
    assert eng.g1_normalize(inf).tobytes() == inf.tobytes()


@pytest.mark.parametrize("n", [256, 257, 300, 1000])
def test_scalar_mul_block_compacted_additions(eng, n):
    """k_jac_mul_compact (a block's additions compacted onto full warps) replays the reference's exact double/add sequence per
    scalar: raw Jacobian bytes equal the oracle's and the thread-per-scalar kernel's (mul_compact = 0), at sizes that leave the
    last block ragged, with scalars 0, 1, 2, r-1, 2^k, all-ones runs, and bases at infinity / with z == 1."""
    import paritytech_bn_b200 as bn
    R = o.R_ORDER
    ks = [0, 1, 2, R - 1, 1 << 200, (1 << 253) - 1, 0x5555555555555555555555555555555555555555555555555555555555555555 % R,
          0xaaaaaaaaaaaaaaaaaaaaaaaaaaaaaaaaaaaaaaaaaaaaaaaaaaaaaaaaaaaaaaaa % R, 3, 1 << 253] + util.scalars(n, stream=17)
    fr = util.fr_bytes(ks[:n])
    b1, b2 = _workload(64, 18)
    idx = np.arange(n) % 64
    g1 = b1.reshape(64, G1)[idx].copy(); g2 = b2.reshape(64, G2)[idx].copy()
    g1[12], g2[12] = util.G1_INF, util.G2_INF
    g1[13], g2[13] = util.G1_GEN, util.G2_GEN
    g1, g2 = g1.reshape(-1), g2.reshape(-1)
    e0, e2 = bn.Engine(0, mul_compact=0), bn.Engine(0, mul_compact=2)      # thread per scalar / compacted for G1 and G2
    try:
        got1, got2 = e2.g1_mul(g1, fr), e2.g2_mul(g2, fr)
        assert got1.tobytes() == c.g1_mul(g1, fr).tobytes() == e0.g1_mul(g1, fr).tobytes() == eng.g1_mul(g1, fr).tobytes()
        assert got2.tobytes() == c.g2_mul(g2, fr).tobytes() == e0.g2_mul(g2, fr).tobytes() == eng.g2_mul(g2, fr).tobytes()
    finally:
        e0.close(); e2.close()


def test_scalar_mul_normalized_vs_oracle(eng):
    """bn254_g{1,2}_mul_normalized_many (signed-window kernel) == the oracle's (p * k).normalize(): edge scalars, extreme
    window digits, z == 1 bases and points at infinity next to random inputs."""
    n = 2048
    ks = util.scalars(n, stream=13)
    base1, base2 = _workload(n, 14)
    R = o.R_ORDER
    sc = [0, 1, 2, R - 1, 8, 9, 16, int("9" * 60, 16) % R, int("f" * 63, 16) % R, int("8" * 63, 16) % R, R - 2] + ks[11:]
    fr = util.fr_bytes(sc)
    base1 = base1.reshape(n, G1).copy(); base2 = base2.reshape(n, G2).copy()
    base1[20], base2[20] = util.G1_GEN, util.G2_GEN
    base1[21], base2[21] = util.G1_INF, util.G2_INF
    base1, base2 = base1.reshape(-1), base2.reshape(-1)
    assert eng.g1_mul_normalized(base1, fr).tobytes() == c.g1_normalize(c.g1_mul(base1, fr)).tobytes()
    n2 = 512
    assert eng.g2_mul_normalized(base2[:n2 * G2], fr[:n2 * FR]).tobytes() == c.g2_normalize(c.g2_mul(base2[:n2 * G2], fr[:n2 * FR])).tobytes()


@pytest.mark.parametrize("which,log2n", [("g1", 20), ("g2", 18)], ids=["g1_mul_n=2^20", "g2_mul_n=2^18"])
def test_scalar_mul_config4_sizes(eng, which, log2n):
    """BASELINE config #4 at its own sizes (2^20 G1 * Fr, 2^18 G2 * Fr).  Bases k_i * G with z != 1 and Montgomery scalars incl.
    0, 1, 2, r-1.  Normalised bytes of the window kernel against the oracle on a 2^16 sample, raw Jacobian bytes of the
    reference-sequence kernel on 2^14, and over ALL elements the two kernels must agree after normalisation."""
    n = 1 << log2n
    unit = G1 if which == "g1" else G2
    gen = util.G1_GEN if which == "g1" else util.G2_GEN
    mul, mul_n, norm = (eng.g1_mul, eng.g1_mul_normalized, eng.g1_normalize) if which == "g1" else (eng.g2_mul, eng.g2_mul_normalized, eng.g2_normalize)
    cmul, cnorm = (c.g1_mul, c.g1_normalize) if which == "g1" else (c.g2_mul, c.g2_normalize)
    base = mul(np.tile(gen, n), util.fr_images(n, stream=91))                        # k_i * G, Jacobian with z != 1
    fr = util.fr_images(n, stream=92)
    fr[:4 * FR] = util.fr_bytes([0, 1, 2, o.R_ORDER - 1])
    raw = mul(base, fr)
    nrm = mul_n(base, fr)
    assert norm(raw).tobytes() == nrm.tobytes()                                      # all n elements, both kernels
    m = 1 << 16
    idx = np.concatenate([np.arange(0, 64), np.arange(64, n, max(1, (n - 64) // (m - 64)))[:m - 64]])
    bs, fs = base.reshape(n, unit)[idx].reshape(-1), fr.reshape(n, FR)[idx].reshape(-1)
    assert nrm.reshape(n, unit)[idx].tobytes() == cnorm(cmul(bs, fs)).tobytes()
    m2 = 1 << 14
    assert raw[:m2 * unit].tobytes() == cmul(base[:m2 * unit], fr[:m2 * FR]).tobytes()  # the reference's own Jacobian representative


def test_group_law_elementwise(eng):
    """bn254_g{1,2}_{add,double,neg,eq}_many: the reference's own Jacobian representatives (raw bytes vs the C oracle for add /
    double, vs the Python oracle for neg and ==), with the zero / equal-point branches of add, infinity, and == across different
    representatives of one point."""
    n = 300
    p, q = _workload(n, 15)
    p2, q2 = _workload(n, 16)
    P, Q, P2, Q2 = (x.reshape(n, -1).copy() for x in (p, q, p2, q2))
    P2[0], Q2[0] = P[0], Q[0]                                # equal points: the doubling branch of add
    P[1], Q[1] = util.G1_INF, util.G2_INF                    # infinity + b = b
    P2[2], Q2[2] = util.G1_INF, util.G2_INF                  # a + infinity = a
    P2[3], Q2[3] = eng.g1_neg(P[3]), eng.g2_neg(Q[3])        # a + (-a) = infinity
    a1, a2, b1, b2 = P.reshape(-1), Q.reshape(-1), P2.reshape(-1), Q2.reshape(-1)
    s1, s2 = eng.g1_add(a1, b1), eng.g2_add(a2, b2)
    assert s1.tobytes() == c.g1_add(a1, b1).tobytes() and s2.tobytes() == c.g2_add(a2, b2).tobytes()
    assert not s1.reshape(n, G1)[3][64:].any() and not s2.reshape(n, G2)[3][128:].any()      # z == 0
    assert eng.g1_double(a1).tobytes() == c.g1_double(a1).tobytes() and eng.g2_double(a2).tobytes() == c.g2_double(a2).tobytes()
    m = 40
    for which, a, unit, F, frm in (("g1", a1, G1, o.FQ_OPS, o.g1_from_bytes), ("g2", a2, G2, o.FQ2_OPS, o.g2_from_bytes)):
        neg = getattr(eng, which + "_neg")(a)
        pts = [frm(a[i * unit:(i + 1) * unit].tobytes()) for i in range(m)]
        assert neg[:m * unit].tobytes() == b"".join(o.to_bytes(o.g_neg(F, x)) for x in pts)
        norm = getattr(eng, which + "_normalize")(a)
        eq = getattr(eng, which + "_eq")
        assert eq(a, norm).all()                                             # same points, other representatives
        assert eq(getattr(eng, which + "_add")(a, a), getattr(eng, which + "_double")(a)).all()
        shifted = np.roll(a.reshape(n, unit), 1, axis=0).reshape(-1)
        want = [int(o.g_eq(F, x, frm(shifted[i * unit:(i + 1) * unit].tobytes()))) for i, x in enumerate(pts)]
        assert list(eq(a, shifted)[:m]) == want and not any(want[2:])
        assert eq(getattr(eng, which + "_add")(a, neg), np.tile(util.G1_INF if which == "g1" else util.G2_INF, n)).all()


def test_gt_algebra_and_joux(eng):
    """Gt::{mul,pow,inverse,==} and the Joux three-party check of examples/joux.rs:7-21."""
    a, b, cc = util.scalars(3, stream=21)
    fa, fb, fc = util.fr_bytes([a]), util.fr_bytes([b]), util.fr_bytes([cc])
    pa, pb, pc_ = (eng.g1_mul(util.G1_GEN, f) for f in (fa, fb, fc))
    qa, qb, qc = (eng.g2_mul(util.G2_GEN, f) for f in (fa, fb, fc))
    e = eng.pairing(np.concatenate([pb, pc_, pa]), np.concatenate([qc, qa, qb]))
    shared = eng.gt_pow(e, np.concatenate([fa, fb, fc]))
    k = shared.reshape(3, GT)
    assert k[0].tobytes() == k[1].tobytes() == k[2].tobytes()
    assert shared.tobytes() == c.gt_pow(e, np.concatenate([fa, fb, fc])).tobytes()
    assert eng.gt_eq(shared, np.tile(k[0], 3)).all()
    assert not eng.gt_eq(e, np.tile(k[0], 3)).any()
    inv, some = eng.gt_inverse(e)
    rinv, rsome = c.fq12_inverse(e)
    assert some.all() and inv.tobytes() == rinv.tobytes()
    assert eng.gt_mul(e, inv).tobytes() == np.tile(util.GT_ONE, 3).tobytes()
    z, some = eng.gt_inverse(np.zeros(GT, np.uint8))
    assert some[0] == 0
    fe, some = eng.final_exponentiation(np.zeros(GT, np.uint8))       # Option::None
    assert some[0] == 0


@pytest.mark.parametrize("log2n", [16], ids=["n_checks=2^16"])
def test_full_size_products_known_answer(eng, log2n):
    """2^16 Groth16-shaped checks (BASELINE config #3 size): valid checks are exactly one, tampered ones are not, and ALL
    Gt values + accept bytes equal the oracle's pairing_batch -- flat 4-pair layout and the shared-G2Precomp entry point.
    Inputs made by the CUDA scalar-mul kernels."""
    n_checks = int(os.environ.get("BN_TEST_FULL_CHECKS", str(1 << log2n)))

    def smul(which, ks):
        fr = util.fr_bytes(ks)
        if which == "g1":
            return eng.g1_mul(np.tile(util.G1_GEN, len(ks)), fr)
        return eng.g2_mul(np.tile(util.G2_GEN, len(ks)), fr)

    w = util.groth16_checks(n_checks, stream=31, scalar_mul=smul)
    out, is_one = eng.pairing_batch(w["g1"], w["g2"], pairs_per_check=4)
    assert (is_one == w["valid"]).all()
    ones = out.reshape(n_checks, GT)[w["valid"] == 1]
    assert (ones == util.GT_ONE).all()
    rout, ris = c.pairing_product_many(w["g1"], w["g2"], 4)
    assert out.tobytes() == rout.tobytes() and (ris == is_one).all()
    pc = eng.g2_precompute(w["vk_g2"])
    out2, is_one2 = eng.pairing_batch_precomp(pc, w["g1_shared"], w["g1_var"], w["g2_var"], 1)
    pc.close()
    assert out2.tobytes() == rout.tobytes() and (is_one2 == is_one).all()


def test_sharded_product(eng):
    """One product split over contexts with the 384-byte partial gather (all on one GPU if only one)."""
    import ctypes

    import paritytech_bn_b200 as bn
    n = 64
    ks1 = util.scalars(n, stream=41)
    ks2 = util.scalars(n, stream=42)
    # sum a_i*b_i == 0 mod r  =>  the product is exactly one (SURVEY 8e known-answer construction)
    tot = sum(x * y for x, y in zip(ks1[:-1], ks2[:-1])) % o.R_ORDER
    ks1[-1] = (-tot * pow(ks2[-1], -1, o.R_ORDER)) % o.R_ORDER
    g1, g2 = util.points_cpu(ks1, ks2)
    import torch
    ndev = min(torch.cuda.device_count(), 4)
    engines = [eng] + [bn.Engine(d) for d in range(1, ndev)]
    if ndev == 1:
        engines.append(bn.Engine(0))      # two contexts on one device still exercise the gather path
    out, is_one = bn.pairing_product_sharded(engines, g1, g2)
    assert is_one and out.tobytes() == util.GT_ONE.tobytes()
    g1[:G1] = util.G1_GEN
    out, is_one = bn.pairing_product_sharded(engines, g1, g2)
    assert not is_one and out.tobytes() == c.pairing_batch(g1, g2).tobytes()
    for e in engines[1:]:
        e.close()


def _engines_for_sharding(eng, want):
    """`want` engines over the visible devices (round-robin): several contexts share a GPU when there are fewer devices."""
    import torch

    import paritytech_bn_b200 as bn
    ndev = torch.cuda.device_count()
    return [eng] + [bn.Engine(d % ndev) for d in range(1, want)]


@pytest.mark.parametrize("log2n", [18], ids=["n_pairs=2^18"])
def test_sharded_product_known_answer_large(eng, log2n):
    """BASELINE config #5's exchange step at size: ONE product of 2^18 pairs split over 4 contexts (4 GPUs when the box has
    them, otherwise contexts sharing devices), 384-byte partials gathered by peer copy, one final exponentiation.
    sum a_i b_i == 0 mod r makes the product exactly one -- no 2^18-pairing CPU run needed; one tampered pair
    must flip the answer; 1-, 2- and 4-way splits give identical bytes; points from the CUDA scalar-mul kernels."""
    import paritytech_bn_b200 as bn
    n = 1 << log2n
    ks1 = util.scalars(n, stream=43)
    ks2 = util.scalars(n, stream=44)
    tot = sum(x * y for x, y in zip(ks1[:-1], ks2[:-1])) % o.R_ORDER
    ks1[-1] = (-tot * pow(ks2[-1], -1, o.R_ORDER)) % o.R_ORDER
    g1 = eng.g1_mul(np.tile(util.G1_GEN, n), util.fr_bytes(ks1))
    g2 = eng.g2_mul(np.tile(util.G2_GEN, n), util.fr_bytes(ks2))
    engines = _engines_for_sharding(eng, 4)
    try:
        out4, one4, ms = bn.pairing_product_sharded(engines, g1, g2, timed=True)
        assert one4 and out4.tobytes() == util.GT_ONE.tobytes()
        assert len(ms) == 3 and all(x >= 0 for x in ms)
        out2, one2 = bn.pairing_product_sharded(engines[:2], g1, g2)
        out1, one1 = bn.pairing_product_sharded(engines[:1], g1, g2)
        assert one2 and one1 and out2.tobytes() == out1.tobytes() == out4.tobytes()
        g1[:G1] = util.G1_GEN                                   # tamper with one pair
        t4, tone4 = bn.pairing_product_sharded(engines, g1, g2)
        t1, tone1 = bn.pairing_product_sharded(engines[:1], g1, g2)
        assert not tone4 and not tone1 and t4.tobytes() == t1.tobytes() != util.GT_ONE.tobytes()
        # every partial is an arbitrary Fq12: the per-context Miller products are not one on their own
        lo, hi = bn.shard_range(n, 1, 4)
        assert eng.miller_product(g1[lo * G1:hi * G1], g2[lo * G2:hi * G2]).tobytes() != util.GT_ONE.tobytes()
    finally:
        for e in engines[1:]:
            e.close()


@pytest.mark.parametrize("n,ways", [(1, 2), (7, 3), (1000, 2), (70001, 4)])
def test_pairing_many_sharded(eng, n, ways):
    """bn254_pairing_many_sharded: ONE call, n independent pairs over a context list -- every result lands in the caller's
    own index range and equals the single-context call (and the oracle on a sample), ragged splits and n < ways included."""
    import paritytech_bn_b200 as bn
    base = min(n, 1024)
    p1, p2 = util.points_cpu(util.scalars(base, stream=45), util.scalars(base, stream=46))
    idx = (np.arange(n) * 2654435761 % base).astype(np.int64)
    g1 = p1.reshape(base, G1)[idx].copy(); g2 = p2.reshape(base, G2)[(idx * 5 + 1) % base].copy()
    if n > 10:
        g1[n // 2] = util.G1_INF
    g1, g2 = g1.reshape(-1), g2.reshape(-1)
    engines = _engines_for_sharding(eng, ways)
    try:
        got = bn.pairing_sharded(engines, g1, g2)
    finally:
        for e in engines[1:]:
            e.close()
    assert got.tobytes() == eng.pairing(g1, g2).tobytes()
    m = min(n, 64)
    sel = np.unique(np.concatenate([np.arange(m), np.arange(n - m, n), [n // 2]]))
    exp = c.pairing_many(g1.reshape(n, G1)[sel].reshape(-1), g2.reshape(n, G2)[sel].reshape(-1))
    assert got.reshape(n, GT)[sel].tobytes() == exp.tobytes()


def test_abi_smoke_compiled_c_consumer():
    """tests/abi/abi_smoke.c, built with gcc against include/bn254_b200.h + libbn254_b200.so alone: struct sizes, argument
    order and NULL-able outputs as a C (or bindgen) user meets them; it runs the reference's test_reduced_pairing vector
    and a 4-pair product."""
    import subprocess

    from tests import abi
    exe = abi.build()
    import paritytech_bn_b200 as bn
    env = dict(os.environ, LD_LIBRARY_PATH=os.path.dirname(bn.library_path()) + ":" + os.environ.get("LD_LIBRARY_PATH", ""))
    p = subprocess.run([exe], capture_output=True, text=True, timeout=600, env=env)
    assert p.returncode == 0, p.stdout + p.stderr
    assert "abi_smoke ok" in p.stdout


@pytest.mark.parametrize("n,quad", [(1, 1), (3, 1), (17, 1), (149, 1), (592, 1), (1, 0), (3, 0), (17, 0), (149, 0), (2368, 0)])
def test_warp_cooperative_latency_path(n, quad):
    """Batches of at most coop_max_n pairings run the generated microcode: four warps per pairing up to coop4_max_n = 592
    (k_coop4_pairing, quad = 1: one to four pairings per block), one warp per pairing above (k_coop_pairing, quad = 0 forces it):
    pairing, miller_loop_many, final_exponentiation (incl. zero -> None), Gt::pow and small pairing_batch products give the
    oracle's bytes, the same bytes as the one-thread-per-pairing kernels (coop_max_n = 0), with points at infinity and z == 1."""
    import paritytech_bn_b200 as bn
    eng = bn.Engine(0) if quad else bn.Engine(0, coop4_max_n=0)
    try:
        _cooperative_case(eng, n)
    finally:
        eng.close()


def _cooperative_case(eng, n):
    import paritytech_bn_b200 as bn
    base = min(n, 256)
    p1, p2 = util.points_cpu(util.scalars(base, stream=91), util.scalars(base, stream=92))
    idx = (np.arange(n) * 2654435761 % base).astype(np.int64)
    g1 = p1.reshape(base, G1)[idx].copy(); g2 = p2.reshape(base, G2)[(idx * 3 + 1) % base].copy()
    if n >= 17:
        g1[3] = util.G1_INF; g2[5] = util.G2_INF; g1[7], g2[7] = util.G1_GEN, util.G2_GEN
    g1, g2 = g1.reshape(-1), g2.reshape(-1)
    exp = c.pairing_many(g1, g2)
    got = eng.pairing(g1, g2)
    assert got.tobytes() == exp.tobytes()
    ml, ok = eng.miller_loop_many(g2, g1)
    rml, rok = c.miller_loop_many(g2, g1)
    assert ml.tobytes() == rml.tobytes() and (ok == rok).all()
    vals = ml.reshape(n, GT).copy()
    if n >= 3:
        vals[1] = 0                                                            # Option::None
    fe, some = eng.final_exponentiation(vals.reshape(-1))
    rfe, rsome = c.final_exp(vals.reshape(-1))
    assert fe.tobytes() == rfe.tobytes() and (some == rsome).all()
    # Gt::pow one warp per element: edge scalars 0, 1, 2, r-1 next to random ones
    ks = ([0, 1, 2, o.R_ORDER - 1] + util.scalars(n, stream=93))[:n]
    pw = eng.gt_pow(got, util.fr_bytes(ks))
    assert pw.tobytes() == c.gt_pow(got, util.fr_bytes(ks)).tobytes()
    if n % 3 == 0:
        out, is_one = eng.pairing_batch(g1, g2, pairs_per_check=3)
        rout, ris = c.pairing_product_many(g1, g2, 3)
        assert out.tobytes() == rout.tobytes() and (is_one == ris).all()
    e0 = bn.Engine(0, coop_max_n=0)                                            # the throughput kernels on the same inputs
    try:
        assert e0.pairing(g1, g2).tobytes() == got.tobytes()
    finally:
        e0.close()


def test_golden_vector_on_the_latency_path(eng, golden):
    """src/groups/mod.rs:928-999 at n = 1 is the warp-cooperative kernel's job; also n = 1 with the path disabled."""
    import paritytech_bn_b200 as bn
    v = golden["test_reduced_pairing"]
    p = eng.g1_mul(util.G1_GEN, util.fr_bytes([int(v["g1_scalar"])]))
    q = eng.g2_mul(util.G2_GEN, util.fr_bytes([int(v["g2_scalar"])]))
    want = util.fq_mont_bytes([int(x) for x in v["expected"]])
    assert eng.pairing(p, q).tobytes() == want
    e0 = bn.Engine(0, coop_max_n=0)
    try:
        assert e0.pairing(p, q).tobytes() == want
    finally:
        e0.close()


def test_shared_memory_ownership_check_build():
    """compute-sanitizer is closed on this pool; its stand-in is the -DBN_SLOT_CHECK build, whose slot accessors verify that
    every shared-memory access of the slot-file and warp-cooperative kernels stays inside the region its issuer owns.
    scripts/slot_check_cases.py runs the small cases of every such kernel on it (results vs the oracle), expects 0 violations,
    and exactly 1 after a deliberate out-of-region store (the checker is live)."""
    import subprocess
    import sys

    import paritytech_bn_b200 as bn
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    if not os.path.exists(bn.library_path("_chk")):
        import __graft_entry__
        __graft_entry__.build()
    p = subprocess.run([sys.executable, os.path.join(root, "scripts", "slot_check_cases.py")], capture_output=True, text=True, timeout=1200)
    assert p.returncode == 0, p.stdout[-3000:] + p.stderr[-3000:]
    assert "slot check ok: 0 violations" in p.stdout


def test_conditional_subtraction_tie_cases(eng):
    """The predicated conditional subtraction decides on the top limb and takes a rare exact path when it ties
    with q's top limb: sums / differences / products landing in [q_top << 224, (q_top + 1) << 224)."""
    rng = np.random.default_rng(77)
    top = (o.Q >> 224) << 224
    sums = [o.Q - 2, o.Q - 1, o.Q, o.Q + 1, o.Q + 2, top, top + 1, (top | ((1 << 224) - 1))]
    sums += [top + int(rng.integers(0, 2**62)) * (1 << 160) + int(rng.integers(0, 2**62)) for _ in range(504)]
    aa, bb = [], []
    for s in sums:
        lo = max(0, s - (o.Q - 1))
        a = lo + (int(rng.integers(0, 2**62)) * (1 << 190)) % (min(s, o.Q - 1) - lo + 1)
        aa.append(a); bb.append(s - a)
    a = np.frombuffer(b"".join(v.to_bytes(32, "little") for v in aa), dtype=np.uint8).copy()   # raw Montgomery images
    b = np.frombuffer(b"".join(v.to_bytes(32, "little") for v in bb), dtype=np.uint8).copy()
    assert eng.field_op("fq_add", a, b).tobytes() == c.fq_add(a, b).tobytes()
    assert eng.field_op("fq_sub", a, b).tobytes() == c.fq_sub(a, b).tobytes()
    assert eng.field_op("fq_sub", b, a).tobytes() == c.fq_sub(b, a).tobytes()
    assert eng.field_op("fq_mul", a, b).tobytes() == c.fq_mul(a, b).tobytes()


def test_mul_by_xi_quotient_estimate_boundaries(eng):
    """fq2_mul_by_xi is two single-pass 9a +- b reductions with a 16-bit quotient estimate (fq_ptx.cuh).  Inputs are
    built so that 9*a1 + a0 and 9*a0 + (q - a1) sit just below / at / above every multiple of q and at the edges of
    the estimate's 2^242 granule; random and extreme elements on top.  Oracle: Fq2 multiplication by the element 9 + u."""
    rng = np.random.default_rng(77)
    Q = o.Q
    pairs = []
    for k in range(0, 11):
        for d in (-2, -1, 0, 1, 2, (1 << 224) - 1, -(1 << 224), 1 << 242, -(1 << 242), (1 << 242) - 1):
            tgt = k * Q + d
            for _ in range(3):
                hi = min(Q - 1, tgt // 9) if tgt >= 0 else -1
                lo = max(0, -(-(tgt - (Q - 1)) // 9))
                if hi < lo:
                    continue
                a9 = int(rng.integers(0, 1 << 62)) % (hi - lo + 1) + lo if hi - lo < (1 << 62) else lo + (util._rand_fq(rng) % (hi - lo + 1))
                bb = tgt - 9 * a9
                if not 0 <= bb < Q:
                    continue
                pairs.append((bb, a9))                        # c1 = 9*a1 + a0 = tgt
                pairs.append((a9, (Q - bb) % Q))              # c0 = 9*a0 - a1 = tgt - q (mod q)
    for v in (0, 1, Q - 1, Q - 2, (Q - 1) // 9, (Q - 1) // 9 + 1, Q // 2):
        for w in (0, 1, Q - 1, Q // 3):
            pairs.append((v, w))
    pairs += [(util._rand_fq(rng), util._rand_fq(rng)) for _ in range(2000)]
    a = np.frombuffer(util.fq_mont_bytes([x for p in pairs for x in p]), dtype=np.uint8).copy()
    xi = np.tile(np.frombuffer(util.fq_mont_bytes([9, 1]), dtype=np.uint8), len(pairs))
    got = eng.field_op("fq2_mul_by_xi", a)
    assert got.tobytes() == c.fq2_mul(a, xi).tobytes()
    exp = b"".join(util.fq_mont_bytes([(9 * x - y) % Q, (9 * y + x) % Q]) for x, y in pairs)
    assert got.tobytes() == exp


def test_fused_and_split_miller_paths_agree(eng):
    """Both Miller implementations (single fused kernel; affine -> line table -> Miller-from-table) give the
    same pre-final-exponentiation bytes as the oracle, including pairs with points at infinity."""
    import paritytech_bn_b200 as bn
    n = 200
    g1, g2 = util.points_cpu(util.scalars(n, stream=61), util.scalars(n, stream=62))
    g1 = g1.reshape(n, G1).copy(); g2 = g2.reshape(n, G2).copy()
    g1[5] = util.G1_INF; g2[9] = util.G2_INF; g1[199] = util.G1_INF; g2[199] = util.G2_INF
    g1, g2 = g1.reshape(-1), g2.reshape(-1)
    exp, exp_ok = c.miller_loop_many(g2, g1)
    got, ok = eng.miller_loop_many(g2, g1)
    assert got.tobytes() == exp.tobytes() and (ok == exp_ok).all()
    e2 = bn.Engine(0, fused_miller=1)
    got2, ok2 = e2.miller_loop_many(g2, g1)
    e2.close()
    assert got2.tobytes() == exp.tobytes() and (ok2 == exp_ok).all()


def test_miller_product_and_distributed_protocol_single_process(eng):
    """bn254_miller_product + the gather/combine protocol of paritytech_bn_b200.distributed, run in one
    process over 3 ragged shards (the gloo world_size-2 test covers the process-group plumbing)."""
    import paritytech_bn_b200 as bn
    n = 11
    g1, g2 = util.points_cpu(util.scalars(n, stream=71), util.scalars(n, stream=72))
    g1 = g1.reshape(n, G1).copy(); g1[4] = util.G1_INF; g1 = g1.reshape(-1)
    assert eng.gt_one().tobytes() == util.GT_ONE.tobytes()
    assert eng.miller_product(g1, g2).tobytes() == c.miller_loop_batch_skip(g2, g1).tobytes()
    parts = []
    for d in range(3):
        lo, hi = bn.shard_range(n, d, 3)
        parts.append(eng.miller_product(g1[lo * G1:hi * G1], g2[lo * G2:hi * G2]))
    out, is_one = bn.combine_partials(eng, np.stack(parts))
    assert out.tobytes() == c.pairing_batch(g1, g2).tobytes() and not is_one
    out1, _ = eng.pairing_batch(g1, g2)
    assert out1.tobytes() == out.tobytes()


@pytest.mark.parametrize("log2n", [16], ids=["n=2^16"])
def test_pairing_full_batch_round_trip_properties(eng, log2n):
    """2^16 pairings (BASELINE config #2 size): ALL outputs byte-compared with the oracle, plus the
    size-independent properties -- bilinearity in the first argument e(2P, Q) == e(P, Q)^2 for every pair, no result one."""
    n = int(os.environ.get("BN_TEST_FULL_PAIRINGS", str(1 << log2n)))
    fr1, fr2 = util.fr_bytes(util.scalars(n, stream=81)), util.fr_bytes(util.scalars(n, stream=82))
    g1 = eng.g1_mul(np.tile(util.G1_GEN, n), fr1)
    g2 = eng.g2_mul(np.tile(util.G2_GEN, n), fr2)
    two = util.fr_bytes([2] * n)
    g1x2 = eng.g1_mul(g1, two)
    e1 = eng.pairing(g1, g2)
    e2 = eng.pairing(g1x2, g2)
    assert e2.tobytes() == eng.gt_mul(e1, e1).tobytes()
    assert not (e1.reshape(n, GT) == util.GT_ONE).all(axis=1).any()
    assert e1.tobytes() == c.pairing_many(g1, g2).tobytes()


def test_pairing_across_table_chunk_boundary(eng):
    """The split Miller path materialises line tables for at most 2^17 pairings at a time: a batch of 2^17 + 37
    pairs crosses the chunk boundary with a ragged tail; indices around the boundary and the tail are compared
    with the oracle bit for bit, and the second chunk must not depend on the first (same pairs, same bytes)."""
    n = (1 << 17) + 37
    base = 4096
    fr1, fr2 = util.fr_bytes(util.scalars(base, stream=85)), util.fr_bytes(util.scalars(base, stream=86))
    p1 = eng.g1_mul(np.tile(util.G1_GEN, base), fr1).reshape(base, G1)
    p2 = eng.g2_mul(np.tile(util.G2_GEN, base), fr2).reshape(base, G2)
    idx = (np.arange(n) * 2654435761 % base).astype(np.int64)       # a fixed pseudo-random arrangement of the 4096 pairs
    g1 = p1[idx].reshape(-1); g2 = p2[(idx * 7 + 3) % base].reshape(-1)
    out = eng.pairing(g1, g2).reshape(n, GT)
    sel = np.concatenate([np.arange(0, 8), np.arange((1 << 17) - 8, (1 << 17) + 8), np.arange(n - 8, n)])
    exp = c.pairing_many(g1.reshape(n, G1)[sel].reshape(-1), g2.reshape(n, G2)[sel].reshape(-1))
    assert out[sel].tobytes() == exp.tobytes()
    # identical pairs give identical bytes wherever they sit in the batch
    key = idx * base + (idx * 7 + 3) % base
    order = np.argsort(key, kind="stable")
    same = key[order][1:] == key[order][:-1]
    a, b = order[1:][same][:2000], order[:-1][same][:2000]
    assert (out[a] == out[b]).all()


@pytest.mark.parametrize("n", [1, 63, 449, 1000, 1600])
def test_slot_file_and_local_memory_paths_agree(eng, n):
    """The slot-file kernels (one 448-thread block per SM with phase barriers: k_miller_table_sync, k_final_exp_hard_sync),
    the 64-thread slot-file hard part and the local-memory kernels produce the oracle's bytes at batch sizes below, at and
    across the block size, including pairs with points at infinity and the raw Miller values."""
    import paritytech_bn_b200 as bn
    g1, g2 = util.points_cpu(util.scalars(n, stream=71), util.scalars(n, stream=72))
    g1 = g1.reshape(n, G1).copy(); g2 = g2.reshape(n, G2).copy()
    if n > 10:
        g1[3] = util.G1_INF; g2[n - 1] = util.G2_INF
    g1, g2 = g1.reshape(-1), g2.reshape(-1)
    m = min(n, 96)
    exp = c.pairing_many(g1[:m * G1], g2[:m * G2], 0)
    exp_f, exp_ok = c.miller_loop_many(g2[:m * G2], g1[:m * G1])
    ref = eng.pairing(g1, g2)
    assert ref[:m * GT].tobytes() == exp.tobytes()
    for opt in (dict(fexp_smem=0, miller_smem=0, lines_smem=0), dict(fexp_smem=1), dict(fexp_smem=2, miller_smem=1),
                dict(big_blocks=2)):    # 512-thread blocks on 6 slots per thread (otherwise chosen from 2^19 pairings up)
        e2 = bn.Engine(0, coop_max_n=0, **opt)
        got = e2.pairing(g1, g2)
        f, ok = e2.miller_loop_many(g2, g1)
        e2.close()
        assert got.tobytes() == ref.tobytes(), opt
        assert f[:m * GT].tobytes() == exp_f.tobytes() and (ok[:m] == exp_ok).all(), opt


def test_environment_switches_still_fill_the_defaults(eng):
    """The BN254_* variables are only the default filler of bn254_ctx_options: an engine created under BN254_FEXP_SMEM=0
    takes the local-memory final exponentiation and still produces the same bytes."""
    import paritytech_bn_b200 as bn
    g1, g2 = util.points_cpu(util.scalars(70, stream=73), util.scalars(70, stream=74))
    os.environ["BN254_FEXP_SMEM"] = "0"
    try:
        e2 = bn.Engine(0)
    finally:
        del os.environ["BN254_FEXP_SMEM"]
    got = e2.pairing(g1, g2)
    e2.close()
    assert got.tobytes() == eng.pairing(g1, g2).tobytes() == c.pairing_many(g1, g2).tobytes()


@pytest.mark.parametrize("v", [1, 4, 8])
def test_product_checks_on_512_thread_blocks(eng, v):
    """k_miller_product_sync on 6 slots per thread (512-thread blocks, big_blocks=2): pairing_batch with v pairs per check and
    pairing_batch_precomp with shared tables, byte-for-byte against the oracle (a 7th-slot access would corrupt a neighbour)."""
    import paritytech_bn_b200 as bn
    n_checks = 1200                                    # > 512, so that sync_block picks 512
    base = 256
    p1, p2 = util.points_cpu(util.scalars(base, stream=75), util.scalars(base, stream=76))
    idx = (np.arange(n_checks * v) * 2654435761 % base).astype(np.int64)
    g1 = p1.reshape(base, G1)[idx].copy(); g2 = p2.reshape(base, G2)[(idx * 3 + 1) % base].copy()
    g1[5] = util.G1_INF
    g1, g2 = g1.reshape(-1), g2.reshape(-1)
    e2 = bn.Engine(0, big_blocks=2)
    try:
        out, is_one = e2.pairing_batch(g1, g2, pairs_per_check=v)
        rout, ris = c.pairing_product_many(g1, g2, v)
        assert out.tobytes() == rout.tobytes() and (is_one == ris).all()
        if v <= 4:
            vk = p2[:3 * G2]
            pc = e2.g2_precompute(vk)
            g1s = p1.reshape(base, G1)[(np.arange(n_checks * 3) * 7 + 2) % base].reshape(-1)
            out2, _ = e2.pairing_batch_precomp(pc, g1s, g1, g2, v)
            pc.close()
            # oracle: the same check as one flat product of 3 + v pairs
            f1 = np.concatenate([g1s.reshape(n_checks, 3 * G1), g1.reshape(n_checks, v * G1)], axis=1).reshape(-1)
            f2 = np.concatenate([np.tile(vk, (n_checks, 1)), g2.reshape(n_checks, v * G2)], axis=1).reshape(-1)
            assert out2.tobytes() == c.pairing_product_many(f1, f2, 3 + v)[0].tobytes()
    finally:
        e2.close()


@pytest.mark.parametrize("name,unit,ref", [
    ("slot_fq6_mul", 192, c.fq6_mul), ("slot_fq12_mul", 384, c.fq12_mul),
])
def test_slot_file_binary_ops(eng, name, unit, ref):
    """The shared-memory slot-file bodies (csrc/bn_slots.cuh: fused Karatsuba stages) one call at a time against the oracle."""
    rng = np.random.default_rng(11)
    a, b = _rand_fq_elems(rng, unit // 32, 300), _rand_fq_elems(rng, unit // 32, 300)
    assert eng.field_op(name, a, b).tobytes() == ref(a, b).tobytes()
    assert eng.field_op(name, a, a).tobytes() == ref(a, a).tobytes()


@pytest.mark.parametrize("name,ref", [("slot_fq12_sqr", c.fq12_sqr), ("slot_cyclotomic_sqr", c.fq12_cyclotomic_squared)])
def test_slot_file_unary_ops(eng, name, ref):
    rng = np.random.default_rng(12)
    a = _rand_fq_elems(rng, 12, 300)
    assert eng.field_op(name, a).tobytes() == ref(a).tobytes()


def test_slot_dot3_bounds(eng):
    """s_dot3 accumulates three Fq2 products unreduced (sums up to 12 q^2) and reduces once per component; its bound argument
    (imaginary part < 1.14 q 2^256 before the conditional subtraction, real part brought into [0, q 2^256)) is exercised with
    every coefficient at q - 1, 0, 1 and mixed extremes next to random inputs."""
    rng = np.random.default_rng(13)
    Q = o.Q
    ext = [Q - 1, 0, 1, Q - 2, (Q - 1) // 2, (Q + 1) // 2]
    rows_a, rows_b = [], []
    for x in ext:
        for y in ext:
            rows_a.append([x] * 6); rows_b.append([y] * 6)                 # all coefficients equal extremes
            rows_a.append([x, 0] * 3); rows_b.append([y, 0] * 3)           # real parts only
            rows_a.append([0, x] * 3); rows_b.append([0, y] * 3)           # imaginary parts only: most negative real sum
            rows_a.append([x, y] * 3); rows_b.append([y, x] * 3)
    def enc(rows):
        return np.frombuffer(b"".join(o.fq_to_mont(v).to_bytes(32, "little") for r in rows for v in r), dtype=np.uint8).copy()
    a = np.concatenate([enc(rows_a), _rand_fq_elems(rng, 6, 2000)])
    b = np.concatenate([enc(rows_b), _rand_fq_elems(rng, 6, 2000)])
    n = len(a) // 192
    got = eng.field_op("slot_dot3", a, b).reshape(n, 192)
    A, B = a.reshape(n, 3, 64), b.reshape(n, 3, 64)
    acc = None
    for k in range(3):
        p = c.fq2_mul(A[:, k].reshape(-1).copy(), B[:, k].reshape(-1).copy())
        if acc is None:
            acc = p
        else:
            acc = np.concatenate([c.fq_add(acc.reshape(n, 2, 32)[:, h].reshape(-1).copy(), p.reshape(n, 2, 32)[:, h].reshape(-1).copy()).reshape(n, 1, 32)
                                  for h in range(2)], axis=1).reshape(-1)
    assert got[:, :64].tobytes() == acc.tobytes()
    assert not got[:, 64:].any()
