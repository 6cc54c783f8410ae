"""Pin the C oracle (oracle/bn254_oracle.c): reference golden vectors + agreement with the Python oracle."""
import random

import numpy as np

from oracle import bn254_ref as o
from oracle import c_oracle as c


def _fq12(dec):
    v = [int(x) for x in dec]
    return (((v[0], v[1]), (v[2], v[3]), (v[4], v[5])), ((v[6], v[7]), (v[8], v[9]), (v[10], v[11])))


def _fq2(dec):
    return (int(dec[0]), int(dec[1]))


def _b(x):
    return np.frombuffer(o.to_bytes(x), dtype=np.uint8)


def test_constants_match_reference_literals(golden):
    m = golden["montgomery_literals"]
    buf = c.constants().tobytes()
    pos = 0

    def take(n):
        nonlocal pos
        out = [int.from_bytes(buf[pos + 32 * i: pos + 32 * i + 32], "little") for i in range(n)]
        pos += 32 * n
        return out

    fq_one, fq_r2, fq_r3, fr_one, fr_r2, fr_r3 = take(6)
    assert [fq_one, fq_r2, fq_r3] == [int(m["fq"][k]) for k in ("one", "rsquared", "rcubed")]
    assert [fr_one, fr_r2, fr_r3] == [int(m["fr"][k]) for k in ("one", "rsquared", "rcubed")]
    inv_q = int.from_bytes(buf[pos:pos + 16], "little")
    inv_r = int.from_bytes(buf[pos + 16:pos + 32], "little")
    pos += 32
    assert inv_q == int(m["fq"]["inv128"]) and inv_r == int(m["fr"]["inv128"])
    assert take(1) == [int(m["fq_non_residue"])]
    assert take(2) == [int(x) for x in m["fq2_nonresidue"]]
    assert take(1) == [int(m["two_inv"])]
    assert take(2) == [int(x) for x in m["g2_coeff_b"]]
    assert take(2) == [int(x) for x in m["twist_mul_by_q_x"]]
    assert take(2) == [int(x) for x in m["twist_mul_by_q_y"]]
    for tab in ("fq6_frobenius_c1", "fq6_frobenius_c2", "fq12_frobenius_c1"):
        for k in ("1", "2", "3"):
            assert take(2) == [int(x) for x in m[tab][k]], (tab, k)
    g1 = take(3)
    assert g1 == [fq_one, int(m["g1_one_y"]), fq_one]
    g2 = take(6)
    assert g2 == [int(x) for x in m["g2_one"]["x"]] + [int(x) for x in m["g2_one"]["y"]] + [fq_one, 0]


def test_fq12_test_vector(golden):
    v = golden["fq12_test_vector"]
    start = _b(_fq12(v["start"]))
    nxt = start
    for _ in range(100):
        nxt = c.fq12_mul(nxt, start)
    cpy = nxt
    for _ in range(10):
        nxt = c.fq12_sqr(nxt)
    for _ in range(10):
        nxt = c.fq12_neg(c.fq12_sub(c.fq12_add(nxt, start), cpy))
    nxt = c.fq12_sqr(nxt)
    assert nxt.tobytes() == o.to_bytes(_fq12(v["finally"]))


def test_cyclotomic_exp(golden):
    v = golden["test_cyclotomic_exp"]
    assert c.fq12_exp_by_neg_z(_b(_fq12(v["orig"]))).tobytes() == o.to_bytes(_fq12(v["expected"]))


def test_prepared_g2_miller_pairing(golden):
    v = golden["test_prepared_g2"]
    g2 = c.g2_mul(_b(o.G2_ONE), np.frombuffer(o.fr_to_bytes(int(v["g2_scalar"])), dtype=np.uint8))
    pc = c.g2_precompute(g2).tobytes()
    exp = o.to_bytes((_fq2(v["q"]["x"]), _fq2(v["q"]["y"])))
    for e in v["coeffs"]:
        exp += o.to_bytes((_fq2(e["ell_0"]), _fq2(e["ell_vw"]), _fq2(e["ell_vv"])))
    assert len(exp) == c.G2PRECOMP and pc == exp

    m = golden["test_miller_loop"]
    g1 = c.g1_mul(_b(o.G1_ONE), np.frombuffer(o.fr_to_bytes(int(m["g1_scalar"])), dtype=np.uint8))
    ml = c.miller_loop_batch(g2, g1)
    assert ml.tobytes() == o.to_bytes(_fq12(m["expected"]))
    mm, ok = c.miller_loop_many(g2, g1)
    assert ok[0] == 1 and mm.tobytes() == ml.tobytes()

    r = golden["test_reduced_pairing"]
    gt = c.pairing_many(g1, g2)
    assert gt.tobytes() == o.to_bytes(_fq12(r["expected"]))
    fe, some = c.final_exp(ml)
    assert some[0] == 1 and fe.tobytes() == gt.tobytes()
    assert c.pairing_batch(g1, g2).tobytes() == gt.tobytes()


def test_random_agreement_with_python_oracle():
    rng = random.Random(0xB254)
    # field ops
    for _ in range(20):
        a, b = rng.randrange(o.Q), rng.randrange(o.Q)
        A, B = _b(a), _b(b)
        assert c.fq_mul(A, B).tobytes() == o.to_bytes((a * b) % o.Q)
        assert c.fq_add(A, B).tobytes() == o.to_bytes((a + b) % o.Q)
        assert c.fq_sub(A, B).tobytes() == o.to_bytes((a - b) % o.Q)
        inv, some = c.fq_inv(A)
        assert some[0] == 1 and inv.tobytes() == o.to_bytes(pow(a, -1, o.Q))
    inv, some = c.fq_inv(_b(0))
    assert some[0] == 0

    def rfq2():
        return (rng.randrange(o.Q), rng.randrange(o.Q))

    def rfq6():
        return (rfq2(), rfq2(), rfq2())

    def rfq12():
        return (rfq6(), rfq6())

    for _ in range(5):
        a, b = rfq12(), rfq12()
        assert c.fq12_mul(_b(a), _b(b)).tobytes() == o.to_bytes(o.fq12_mul(a, b))
        assert c.fq12_sqr(_b(a)).tobytes() == o.to_bytes(o.fq12_sqr(a))
        ia, some = c.fq12_inverse(_b(a))
        assert some[0] == 1 and ia.tobytes() == o.to_bytes(o.fq12_inv(a))
        for p in (1, 2, 3):
            assert c.fq12_frobenius(_b(a), p).tobytes() == o.to_bytes(o.fq12_frobenius(a, p))
        ell = (rfq2(), rfq2(), rfq2())
        assert c.fq12_mul_by_024(_b(a), _b(ell)).tobytes() == o.to_bytes(o.fq12_mul_by_024(a, *ell))
        assert c.fq6_mul(_b(a[0]), _b(b[0])).tobytes() == o.to_bytes(o.fq6_mul(a[0], b[0]))
        k = rng.randrange(o.R_ORDER)
        kb = np.frombuffer(o.fr_to_bytes(k), dtype=np.uint8)
        assert c.gt_pow(_b(a), kb).tobytes() == o.to_bytes(o.fq12_pow(a, k))
    # groups: raw Jacobian outputs must match too (same add/double sequence)
    for _ in range(4):
        k1, k2, s = rng.randrange(1, o.R_ORDER), rng.randrange(1, o.R_ORDER), rng.randrange(o.R_ORDER)
        p = o.g_mul(o.FQ_OPS, o.G1_ONE, k1)
        q = o.g_mul(o.FQ2_OPS, o.G2_ONE, k2)
        sb = np.frombuffer(o.fr_to_bytes(s), dtype=np.uint8)
        assert c.g1_mul(_b(p), sb).tobytes() == o.to_bytes(o.g_mul(o.FQ_OPS, p, s))
        assert c.g2_mul(_b(q), sb).tobytes() == o.to_bytes(o.g_mul(o.FQ2_OPS, q, s))
        assert c.g1_normalize(_b(p)).tobytes() == o.to_bytes(o.g_normalize(o.FQ_OPS, p))
        assert c.g2_normalize(_b(q)).tobytes() == o.to_bytes(o.g_normalize(o.FQ2_OPS, q))
        assert c.pairing_many(_b(p), _b(q)).tobytes() == o.to_bytes(o.pairing(p, q))


def test_edge_semantics():
    z1, z2 = o.g_zero(o.FQ_OPS), o.g_zero(o.FQ2_OPS)
    p = o.g_mul(o.FQ_OPS, o.G1_ONE, 7)
    q = o.g_mul(o.FQ2_OPS, o.G2_ONE, 11)
    one = o.to_bytes(o.FQ12_ONE)
    assert c.pairing_many(_b(z1), _b(q)).tobytes() == one
    assert c.pairing_many(_b(p), _b(z2)).tobytes() == one
    assert c.pairing_batch(np.zeros(0, np.uint8), np.zeros(0, np.uint8)).tobytes() == one
    # skip semantics: (p,q),(inf,q),(p,inf) == pairing(p,q)
    ps = np.concatenate([_b(p), _b(z1), _b(p)])
    qs = np.concatenate([_b(q), _b(q), _b(z2)])
    assert c.pairing_batch(ps, qs).tobytes() == c.pairing_many(_b(p), _b(q)).tobytes()
    assert c.miller_loop_batch(_b(z2), _b(p)) is None
    assert c.miller_loop_batch(_b(q), _b(z1)) is None
    # product of e(aP, bQ) e(-abP, Q) == 1
    a, b = 123456789, 987654321
    pa = o.g_mul(o.FQ_OPS, o.G1_ONE, a)
    qb = o.g_mul(o.FQ2_OPS, o.G2_ONE, b)
    pn = o.g_neg(o.FQ_OPS, o.g_mul(o.FQ_OPS, o.G1_ONE, a * b % o.R_ORDER))
    out, is_one = c.pairing_product_many(np.concatenate([_b(pa), _b(pn)]), np.concatenate([_b(qb), _b(o.G2_ONE)]), 2)
    assert is_one[0] == 1 and out.tobytes() == one
    # scalar edge cases 0, 1, r-1
    for k in (0, 1, 2, o.R_ORDER - 1):
        kb = np.frombuffer(o.fr_to_bytes(k), dtype=np.uint8)
        assert c.g1_mul(_b(p), kb).tobytes() == o.to_bytes(o.g_mul(o.FQ_OPS, p, k))
        assert c.g2_mul(_b(q), kb).tobytes() == o.to_bytes(o.g_mul(o.FQ2_OPS, q, k))
