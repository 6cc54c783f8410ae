"""Pin the Python oracle against the reference crate's own known-answer vectors.

Fixture: tests/golden/reference_vectors.json (made by tests/golden/extract_golden.py from
/root/reference; provenance file:line is stored inside the fixture).
"""
from oracle import bn254_ref as o


def _fq12(dec):
    c = [int(x) for x in dec]
    return (((c[0], c[1]), (c[2], c[3]), (c[4], c[5])), ((c[6], c[7]), (c[8], c[9]), (c[10], c[11])))


def _fq2(dec):
    return (int(dec[0]), int(dec[1]))


def test_str_vectors(golden):
    # src/fields/mod.rs:67-81
    assert int(golden["test_str"]["fr_minus_one"]) == o.R_ORDER - 1
    assert int(golden["test_str"]["fq_minus_one"]) == o.Q - 1


def test_naf_table(golden):
    assert golden["ate_loop_count_naf"]["digits"] == o.ATE_LOOP_COUNT_NAF


def test_montgomery_literals(golden):
    """Constants the oracle derives from q, r, xi equal the reference's inline Montgomery literals."""
    m = golden["montgomery_literals"]
    fm = o.fq_from_mont
    assert int(m["fq"]["modulus"]) == o.Q and int(m["fr"]["modulus"]) == o.R_ORDER
    assert int(m["fq"]["one"]) == o.MONT_R % o.Q and int(m["fr"]["one"]) == o.MONT_R % o.R_ORDER
    assert int(m["fq"]["rsquared"]) == pow(o.MONT_R, 2, o.Q) and int(m["fr"]["rsquared"]) == pow(o.MONT_R, 2, o.R_ORDER)
    assert int(m["fq"]["rcubed"]) == pow(o.MONT_R, 3, o.Q) and int(m["fr"]["rcubed"]) == pow(o.MONT_R, 3, o.R_ORDER)
    assert int(m["fq"]["inv128"]) == (-pow(o.Q, -1, 1 << 128)) % (1 << 128)
    assert int(m["fr"]["inv128"]) == (-pow(o.R_ORDER, -1, 1 << 128)) % (1 << 128)
    assert int(m["bn_u"]) == o.BN_U
    assert fm(int(m["fq_non_residue"])) == o.Q - 1
    assert tuple(fm(int(x)) for x in m["fq2_nonresidue"]) == o.XI
    assert fm(int(m["two_inv"])) == o.TWO_INV
    assert fm(int(m["g1_one_y"])) == o.G1_ONE[1] and fm(int(m["g1_coeff_b"])) == o.G1_B
    assert tuple(fm(int(x)) for x in m["g2_one"]["x"]) == o.G2_ONE[0]
    assert tuple(fm(int(x)) for x in m["g2_one"]["y"]) == o.G2_ONE[1]
    assert tuple(fm(int(x)) for x in m["g2_coeff_b"]) == o.G2_B
    assert tuple(fm(int(x)) for x in m["twist_mul_by_q_x"]) == o.TWIST_MUL_BY_Q_X
    assert tuple(fm(int(x)) for x in m["twist_mul_by_q_y"]) == o.TWIST_MUL_BY_Q_Y
    for k in ("1", "2", "3"):
        assert tuple(fm(int(x)) for x in m["fq6_frobenius_c1"][k]) == o.FQ6_FROB_C1[int(k)]
        assert tuple(fm(int(x)) for x in m["fq6_frobenius_c2"][k]) == o.FQ6_FROB_C2[int(k)]
        assert tuple(fm(int(x)) for x in m["fq12_frobenius_c1"][k]) == o.FQ12_FROB_C1[int(k)]


def test_generators_on_curve(golden):
    assert o.affine_on_curve(o.FQ_OPS, o.G1_B, o.G1_ONE[0], o.G1_ONE[1])
    assert o.affine_on_curve(o.FQ2_OPS, o.G2_B, o.G2_ONE[0], o.G2_ONE[1])
    pp = golden["predefined_pair"]
    assert (int(pp["g1"][0]), int(pp["g1"][1])) == o.G1_ONE[:2]
    assert _fq2(pp["g2_x"]) == o.G2_ONE[0] and _fq2(pp["g2_y"]) == o.G2_ONE[1]


def test_fq12_test_vector(golden):
    # src/fields/mod.rs:93-227
    v = golden["fq12_test_vector"]
    start = _fq12(v["start"])
    nxt = start
    for _ in range(100):
        nxt = o.fq12_mul(nxt, start)
    cpy = nxt
    for _ in range(10):
        nxt = o.fq12_sqr(nxt)
    for _ in range(10):
        nxt = o.fq12_add(nxt, start)
        nxt = o.fq12_sub(nxt, cpy)
        nxt = o.fq12_neg(nxt)
    nxt = o.fq12_sqr(nxt)
    assert nxt == _fq12(v["finally"])


def test_cyclotomic_exp(golden):
    # src/fields/mod.rs:229-344
    v = golden["test_cyclotomic_exp"]
    assert o.fq12_exp_by_neg_z(_fq12(v["orig"])) == _fq12(v["expected"])


def test_prepared_g2(golden):
    # src/groups/mod.rs:779-892: G2*Fr, to_affine and all 87 EllCoeffs
    v = golden["test_prepared_g2"]
    g2 = o.g_mul(o.FQ2_OPS, o.G2_ONE, int(v["g2_scalar"]))
    qa = o.g_to_affine(o.FQ2_OPS, g2)
    assert qa == (_fq2(v["q"]["x"]), _fq2(v["q"]["y"]))
    q, coeffs = o.g2_precompute(qa)
    assert len(coeffs) == 87
    for got, exp in zip(coeffs, v["coeffs"]):
        assert got == (_fq2(exp["ell_0"]), _fq2(exp["ell_vw"]), _fq2(exp["ell_vv"]))


def test_miller_loop(golden):
    # src/groups/mod.rs:642-691: the pre-final-exponentiation value
    v = golden["test_miller_loop"]
    g1 = o.g_mul(o.FQ_OPS, o.G1_ONE, int(v["g1_scalar"]))
    g2 = o.g_mul(o.FQ2_OPS, o.G2_ONE, int(v["g2_scalar"]))
    f = o.miller_loop(o.g2_precompute(o.g_to_affine(o.FQ2_OPS, g2)), o.g_to_affine(o.FQ_OPS, g1))
    assert f == _fq12(v["expected"])


def test_reduced_pairing(golden):
    # src/groups/mod.rs:928-999
    v = golden["test_reduced_pairing"]
    g1 = o.g_mul(o.FQ_OPS, o.G1_ONE, int(v["g1_scalar"]))
    g2 = o.g_mul(o.FQ2_OPS, o.G2_ONE, int(v["g2_scalar"]))
    assert o.pairing(g1, g2) == _fq12(v["expected"])


def test_batch_equals_product_and_edges():
    """pairing_batch semantics (src/groups/mod.rs:904-926): skip infinity, empty -> one, product identity."""
    p1 = o.g_mul(o.FQ_OPS, o.G1_ONE, 5)
    p2 = o.g_mul(o.FQ_OPS, o.G1_ONE, 77)
    q1 = o.g_mul(o.FQ2_OPS, o.G2_ONE, 9)
    q2 = o.g_mul(o.FQ2_OPS, o.G2_ONE, 1234567)
    assert o.pairing_batch([], []) == o.FQ12_ONE
    assert o.pairing(o.g_zero(o.FQ_OPS), q1) == o.FQ12_ONE
    assert o.pairing(p1, o.g_zero(o.FQ2_OPS)) == o.FQ12_ONE
    b = o.pairing_batch([p1, o.g_zero(o.FQ_OPS), p2], [q1, q2, q2])
    assert b == o.fq12_mul(o.pairing(p1, q1), o.pairing(p2, q2))
    # bilinearity: e(5P, 9Q) == e(P,Q)^45
    e = o.pairing(o.G1_ONE, o.G2_ONE)
    assert o.pairing(p1, q1) == o.fq12_pow(e, 45)
    # lib::miller_loop_batch error path
    assert o.lib_miller_loop_batch([(o.g_zero(o.FQ2_OPS), p1)]) is None


def test_g2_subgroup_and_group_law():
    assert o.g2_in_subgroup(*o.G2_ONE[:2])
    a = o.g_mul(o.FQ_OPS, o.G1_ONE, 12345)
    b = o.g_mul(o.FQ_OPS, o.G1_ONE, 54321)
    assert o.g_eq(o.FQ_OPS, o.g_add(o.FQ_OPS, a, b), o.g_mul(o.FQ_OPS, o.G1_ONE, 66666))
    assert o.g_is_zero(o.FQ_OPS, o.g_add(o.FQ_OPS, o.g_mul(o.FQ_OPS, o.G1_ONE, o.R_ORDER - 1), o.G1_ONE))


def test_wnaf3_recoding_of_u_in_device_header():
    """The width-3 signed-window digits of u used by fq12_exp_by_neg_z (bn_tower.cuh) reconstruct u."""
    import os
    import re
    src = open(os.path.join(os.path.dirname(__file__), "..", "paritytech-bn_b200", "csrc", "bn_tower.cuh")).read()
    m = {k: int(re.search(r"#define BN_U_WNAF3_%s (0x[0-9a-f]+)ull" % k, src).group(1), 16) for k in ("NZ", "NEG", "M3")}
    v = 0
    for i in range(64):
        if (m["NZ"] >> i) & 1:
            d = 3 if (m["M3"] >> i) & 1 else 1
            v += (-d if (m["NEG"] >> i) & 1 else d) << i
    assert v == 4965661367192848881                      # src/fields/fq12.rs:122
    assert m["NEG"] & ~m["NZ"] == 0 and m["M3"] & ~m["NZ"] == 0 and (m["NZ"] >> 62) == 1 and not (m["NEG"] >> 62) & 1


def test_wnaf4_recoding_of_u_in_device_header():
    """The width-4 digits (default window of fq12_exp_by_neg_z): magnitude 2 * (I1 I0) + 1, 14 non-zero digits."""
    import os
    import re
    src = open(os.path.join(os.path.dirname(__file__), "..", "paritytech-bn_b200", "csrc", "bn_tower.cuh")).read()
    m = {k: int(re.search(r"#define BN_U_WNAF4_%s (0x[0-9a-f]+)ull" % k, src).group(1), 16) for k in ("NZ", "NEG", "I0", "I1")}
    v, nz = 0, 0
    for i in range(64):
        if (m["NZ"] >> i) & 1:
            d = 2 * (((m["I0"] >> i) & 1) | (((m["I1"] >> i) & 1) << 1)) + 1
            v += (-d if (m["NEG"] >> i) & 1 else d) << i
            nz += 1
    assert v == 4965661367192848881 and nz == 14          # src/fields/fq12.rs:122
    assert (m["NEG"] | m["I0"] | m["I1"]) & ~m["NZ"] == 0
    assert (m["NZ"] >> 62) == 1 and not ((m["NEG"] | m["I0"] | m["I1"]) >> 62) & 1   # top digit is +1
    assert re.search(r"#define BN_U_WINDOW 4", src)


def test_doubling_mask_matches_naf():
    """BN_DBL_MASK_* in bn_curve.cuh marks exactly the doubling-step coefficients of precompute()'s 87 triples."""
    import os
    import re
    from oracle import bn254_ref as o
    src = open(os.path.join(os.path.dirname(__file__), "..", "paritytech-bn_b200", "csrc", "bn_curve.cuh")).read()
    lo = int(re.search(r"#define BN_DBL_MASK_LO (0x[0-9a-f]+)ull", src).group(1), 16)
    hi = int(re.search(r"#define BN_DBL_MASK_HI (0x[0-9a-f]+)ull", src).group(1), 16)
    bits = []
    for d in o.ATE_LOOP_COUNT_NAF:
        bits.append(1)
        if d != 0:
            bits.append(0)
    bits += [0, 0]
    assert len(bits) == 87
    assert lo | (hi << 64) == sum(b << i for i, b in enumerate(bits))


def test_bgcd_model_and_device_constants():
    import os
    """tools/bgcd_model.py (the Python model of fq_inv_bgcd, every bound of the C code asserted inside) converges to (0, 1) with
    the inverse on extremes and random residues; the constants in csrc/bn_tower.cuh are the model's."""
    import re
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    sys.path.insert(0, os.path.join(root, "tools"))
    import bgcd_model as m
    n, bad = m.self_test(400, seed=9)
    assert n > 400 and bad == 0
    src = open(os.path.join(root, "paritytech-bn_b200", "csrc", "bn_tower.cuh")).read()
    assert int(re.search(r"#define BN_BGCD_K (\d+)", src).group(1)) == m.K
    assert int(re.search(r"#define BN_BGCD_ROUNDS (\d+)", src).group(1)) == m.ROUNDS
    assert int(re.search(r"#define BN_BGCD_NEGQINV 0x([0-9a-f]+)u", src).group(1), 16) == m.NEGQINV
    assert m.K * m.ROUNDS >= 2 * 254 - 1


def test_ptx_generator_simulates_and_committed_header_is_current(tmp_path):
    """tools/gen_fq_ptx.py: every generated carry-chain routine (multiplication, wide product, both reductions, add / sub, 9a +- b,
    the 9-limb reduction of the lazy sums, the wide add / sub pair and the borrow-free subtraction) is SIMULATED against Python
    integers before it is printed; the header in the tree is exactly what the generator prints."""
    import os
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    out = tmp_path / "fq_ptx.cuh"
    p = subprocess.run([sys.executable, os.path.join(root, "tools", "gen_fq_ptx.py"), "--out", str(out)], capture_output=True, text=True, timeout=600)
    assert p.returncode == 0, p.stdout[-2000:] + p.stderr[-2000:]
    assert "simulated OK" in p.stdout and "9-limb reductions" in p.stdout
    assert out.read_text() == open(os.path.join(root, "paritytech-bn_b200", "csrc", "fq_ptx.cuh")).read()
