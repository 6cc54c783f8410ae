"""The device headers (paritytech-bn_b200/csrc/*.cuh), compiled for the host with the portable limb
back end, against the C oracle: every formula, constant and branch of the CUDA code path is checked
on the CPU; the GPU run then only has to confirm the PTX primitives and the kernels' data movement."""
import ctypes
import random

import numpy as np
import pytest

from oracle import bn254_ref as o
from oracle import c_oracle as c
from tests import hostsim

FQ, FQ2, FQ6, FQ12, G1, G2, ELL = 32, 64, 192, 384, 96, 192, 192
rng = random.Random(20250)


def _p(a):
    return a.ctypes.data_as(ctypes.c_void_p)


def _b(x):
    return np.frombuffer(o.to_bytes(x), dtype=np.uint8).copy()


def rfq():
    return rng.choice([0, 1, o.Q - 1, rng.randrange(o.Q), rng.randrange(o.Q), rng.randrange(o.Q)])


def rand_elems(n_fq, n):
    return np.concatenate([_b(tuple(rfq() for _ in range(n_fq))) for _ in range(n)])


def hs_bin(name, a, b, unit):
    n = a.size // unit
    out = np.empty_like(a)
    getattr(hostsim.lib(), "hs_" + name)(_p(a), _p(b), _p(out), ctypes.c_size_t(n))
    return out


def hs_un(name, a, unit):
    n = a.size // unit
    out = np.empty_like(a)
    getattr(hostsim.lib(), "hs_" + name)(_p(a), _p(out), ctypes.c_size_t(n))
    return out


@pytest.mark.parametrize("name,unit,ref", [
    ("fq_mul", FQ, c.fq_mul), ("fq_add", FQ, c.fq_add), ("fq_sub", FQ, c.fq_sub),
    ("fq2_mul", FQ2, c.fq2_mul), ("fq6_mul", FQ6, c.fq6_mul), ("fq12_mul", FQ12, c.fq12_mul),
    ("fq2_mul_alias", FQ2, c.fq2_mul), ("fq6_mul_alias", FQ6, c.fq6_mul), ("fq12_mul_alias", FQ12, c.fq12_mul),
])
def test_binary_field_ops(name, unit, ref):
    n = 64
    a, b = rand_elems(unit // 32, n), rand_elems(unit // 32, n)
    assert hs_bin(name, a, b, unit).tobytes() == ref(a, b).tobytes()


@pytest.mark.parametrize("name,unit,ref", [
    ("fq2_sqr", FQ2, c.fq2_sqr), ("fq12_sqr", FQ12, c.fq12_sqr),
    ("fq6_sqr", FQ6, lambda a: c.fq6_mul(a, a)),
    ("fq2_mul_by_xi", FQ2, lambda a: c.fq2_mul(a, np.tile(_b(o.XI), a.size // FQ2))),
    ("fq2_half", FQ2, lambda a: c.fq2_mul(a, np.tile(_b((o.TWO_INV, 0)), a.size // FQ2))),
])
def test_unary_field_ops(name, unit, ref):
    a = rand_elems(unit // 32, 64)
    assert hs_un(name, a, unit).tobytes() == ref(a).tobytes()


def test_inversions_and_frobenius():
    a = rand_elems(1, 40)
    out = np.empty_like(a)
    some = np.empty(40, dtype=np.uint8)
    ro, rs = c.fq_inv(a)
    for fn in ("hs_fq_inv", "hs_fq_inv_w4", "hs_fq_inv_eea"):     # Fermat ladder, its 4-bit window, the uniform binary extended Euclid
        getattr(hostsim.lib(), fn)(_p(a), _p(out), _p(some), ctypes.c_size_t(40))
        assert (some == rs).all() and out.tobytes() == ro.tobytes()
    # the Euclid loop at its extremes: 1, 2, q - 1, q - 2, powers of two (long runs of halvings), values next to q / 2, the
    # Montgomery one and R^2; plus 2,000 random residues
    vals = [1, 2, 3, o.Q - 1, o.Q - 2, (o.Q - 1) // 2, (o.Q + 1) // 2, o.MONT_R % o.Q, (o.MONT_R * o.MONT_R) % o.Q, 0]
    vals += [(1 << k) % o.Q for k in (1, 31, 32, 63, 64, 127, 128, 253, 254, 255)] + [o.Q - (1 << k) for k in (1, 32, 200, 253)]
    vals += [rng.randrange(o.Q) for _ in range(2000)]
    a = np.frombuffer(b"".join(v.to_bytes(32, "little") for v in vals), dtype=np.uint8).copy()      # raw Montgomery images
    out, some = np.empty_like(a), np.empty(len(vals), dtype=np.uint8)
    hostsim.lib().hs_fq_inv_eea(_p(a), _p(out), _p(some), ctypes.c_size_t(len(vals)))
    for i, v in enumerate(vals):
        got = int.from_bytes(out[32 * i:32 * i + 32].tobytes(), "little")
        if v == 0:
            assert some[i] == 0 and got == 0
        else:       # image v = a R  ->  image of the inverse = a^-1 R = R^2 / v
            assert some[i] == 1 and got == pow(v, -1, o.Q) * o.MONT_R * o.MONT_R % o.Q, i
    # the same inverse in 17 rounds of 30 approximated steps (fq_inv_bgcd): status 1 everywhere, i.e. the plain loop is never needed
    vals += [rng.randrange(1, 1 << rng.randrange(1, 254)) for _ in range(3000)] + [o.Q - rng.randrange(1, 1 << rng.randrange(1, 250)) for _ in range(3000)]
    vals += [(1 << k) + d for k in range(1, 254) for d in (-1, 0, 1) if 0 < (1 << k) + d < o.Q]
    vals += [rng.randrange(o.Q) for _ in range(20000)]
    a = np.frombuffer(b"".join(v.to_bytes(32, "little") for v in vals), dtype=np.uint8).copy()
    out, some = np.empty_like(a), np.empty(len(vals), dtype=np.uint8)
    hostsim.lib().hs_fq_inv_bgcd(_p(a), _p(out), _p(some), ctypes.c_size_t(len(vals)))
    r2 = o.MONT_R * o.MONT_R % o.Q
    for i, v in enumerate(vals):
        got = int.from_bytes(out[32 * i:32 * i + 32].tobytes(), "little")
        if v == 0:
            assert some[i] == 0 and got == 0
        else:
            assert some[i] == 1 and got == pow(v, -1, o.Q) * r2 % o.Q, (i, hex(v), some[i])
    a = np.concatenate([rand_elems(12, 6), _b(o.FQ12_ZERO)])
    out = np.empty_like(a)
    some = np.empty(7, dtype=np.uint8)
    hostsim.lib().hs_fq12_inv(_p(a), _p(out), _p(some), ctypes.c_size_t(7))
    ro, rs = c.fq12_inverse(a)
    assert (some == rs).all() and some[6] == 0 and out.tobytes() == ro.tobytes()
    for power in (1, 2, 3):
        out = np.empty_like(a)
        hostsim.lib().hs_fq12_frobenius(_p(a), ctypes.c_int(power), _p(out), ctypes.c_size_t(7))
        assert out.tobytes() == c.fq12_frobenius(a, power).tobytes()


def test_sparse_mul_and_cyclotomic():
    f = rand_elems(12, 16)
    ell = rand_elems(6, 16)
    out = np.empty_like(f)
    hostsim.lib().hs_fq12_mul_by_024(_p(f), _p(ell), _p(out), ctypes.c_size_t(16))
    assert out.tobytes() == c.fq12_mul_by_024(f, ell).tobytes()
    # cyclotomic-subgroup elements: easy part of the final exponentiation of random elements
    g = rand_elems(12, 4)
    cyc = np.concatenate([_b(o.fq12_final_exp_first_chunk(o.fq12_from_bytes(g[i * 384:(i + 1) * 384].tobytes()))) for i in range(4)])
    assert hs_un("fq12_cyclotomic_sqr", cyc, FQ12).tobytes() == c.fq12_cyclotomic_squared(cyc).tobytes()
    assert hs_un("fq12_exp_by_neg_z", cyc, FQ12).tobytes() == c.fq12_exp_by_neg_z(cyc).tobytes()
    out = np.empty_like(g)
    some = np.empty(4, dtype=np.uint8)
    hostsim.lib().hs_final_exp(_p(g), _p(out), _p(some), ctypes.c_size_t(4))
    ro, rs = c.final_exp(g)
    assert (some == rs).all() and out.tobytes() == ro.tobytes()


def _points(n, edge=True):
    ks = [rng.randrange(1, o.R_ORDER) for _ in range(n)]
    g1 = [o.g_mul(o.FQ_OPS, o.G1_ONE, k) for k in ks]
    g2 = [o.g_mul(o.FQ2_OPS, o.G2_ONE, k + 1) for k in ks]
    if edge:
        g1 += [o.g_zero(o.FQ_OPS), o.G1_ONE]
        g2 += [o.G2_ONE, o.g_zero(o.FQ2_OPS)]
    return np.concatenate([_b(p) for p in g1]), np.concatenate([_b(p) for p in g2])


def test_groups_and_scalar_mul():
    p, q = _points(5)
    n = p.size // G1
    p2 = np.roll(p.reshape(n, G1), 1, axis=0).reshape(-1).copy()
    q2 = np.roll(q.reshape(n, G2), 1, axis=0).reshape(-1).copy()
    assert hs_bin("g1_add", p, p2, G1).tobytes() == c.g1_add(p, p2).tobytes()
    assert hs_bin("g2_add", q, q2, G2).tobytes() == c.g2_add(q, q2).tobytes()
    assert hs_bin("g1_add", p, p, G1).tobytes() == c.g1_add(p, p).tobytes()      # equal points -> double branch
    assert hs_un("g1_double", p, G1).tobytes() == c.g1_double(p).tobytes()
    assert hs_un("g2_double", q, G2).tobytes() == c.g2_double(q).tobytes()
    assert hs_un("g1_normalize", p, G1).tobytes() == c.g1_normalize(p).tobytes()
    assert hs_un("g2_normalize", q, G2).tobytes() == c.g2_normalize(q).tobytes()
    ks = [0, 1, 2, o.R_ORDER - 1] + [rng.randrange(o.R_ORDER) for _ in range(n - 4)]
    fr = np.concatenate([np.frombuffer(o.fr_to_bytes(k), dtype=np.uint8) for k in ks]).copy()
    can = np.empty_like(fr)
    hostsim.lib().hs_fr_to_canonical(_p(fr), _p(can), ctypes.c_size_t(n))
    assert can.tobytes() == b"".join(k.to_bytes(32, "little") for k in ks)
    out = np.empty_like(p)
    hostsim.lib().hs_g1_mul(_p(p), _p(fr), _p(out), ctypes.c_size_t(n))
    assert out.tobytes() == c.g1_mul(p, fr).tobytes()
    out = np.empty_like(q)
    hostsim.lib().hs_g2_mul(_p(q), _p(fr), _p(out), ctypes.c_size_t(n))
    assert out.tobytes() == c.g2_mul(q, fr).tobytes()


def _group_cases():
    """(a, b, expected ==) over G1 and G2: equal points in different Jacobian representatives, different points, negatives,
    infinity against infinity (any (x, y, 0)) and against a finite point."""
    p, q = _points(6)
    n = p.size // G1
    P, Q = p.reshape(n, G1).copy(), q.reshape(n, G2).copy()
    pn, qn = c.g1_normalize(p).reshape(n, G1), c.g2_normalize(q).reshape(n, G2)
    inf1 = np.frombuffer(o.to_bytes((0, 1, 0)), dtype=np.uint8)
    inf1b = np.frombuffer(o.to_bytes((5, 7, 0)), dtype=np.uint8)
    inf2 = np.frombuffer(o.to_bytes(((0, 0), (1, 0), (0, 0))), dtype=np.uint8)
    a1 = np.stack([P[0], P[1], P[2], inf1, inf1, P[3]])
    b1 = np.stack([pn[0], P[2], pn[2], inf1b, P[4], inf1])
    a2 = np.stack([Q[0], Q[1], Q[2], inf2, inf2, Q[3]])
    b2 = np.stack([qn[0], Q[2], qn[2], inf2, Q[4], inf2])
    return a1.reshape(-1), b1.reshape(-1), a2.reshape(-1), b2.reshape(-1), [1, 0, 1, 1, 0, 0]


def test_group_neg_and_projective_equality_host_build():
    """jac_neg / jac_eq (src/groups/mod.rs:336-350, :169-195) of the device headers against the Python oracle."""
    a1, b1, a2, b2, want = _group_cases()
    n = len(want)
    for name, a, b, unit, F, frm in (("g1", a1, b1, G1, o.FQ_OPS, o.g1_from_bytes), ("g2", a2, b2, G2, o.FQ2_OPS, o.g2_from_bytes)):
        eq = np.empty(n, dtype=np.uint8)
        getattr(hostsim.lib(), "hs_%s_eq" % name)(_p(a), _p(b), _p(eq), ctypes.c_size_t(n))
        pts_a = [frm(a[i * unit:(i + 1) * unit].tobytes()) for i in range(n)]
        pts_b = [frm(b[i * unit:(i + 1) * unit].tobytes()) for i in range(n)]
        assert list(eq) == [int(o.g_eq(F, x, y)) for x, y in zip(pts_a, pts_b)] == want
        neg = hs_un("%s_neg" % name, a, unit)
        assert neg.tobytes() == b"".join(o.to_bytes(o.g_neg(F, x)) for x in pts_a)
        # p + (-p) is the point at infinity, and == says so
        add = hs_bin("%s_add" % name, a, neg, unit)
        inf = np.tile(np.frombuffer(o.to_bytes(o.g_zero(F)), dtype=np.uint8), n)
        getattr(hostsim.lib(), "hs_%s_eq" % name)(_p(add), _p(inf), _p(eq), ctypes.c_size_t(n))
        assert eq.all()


def test_window_scalar_mul_normalized_host_build():
    """jac_mul_normalized (signed 4-bit window, the body of k_jac_mul_window) == the oracle's (p * k).normalize(), bytes and
    all: random scalars, the edge scalars 0 / 1 / 2 / r-1, scalars made of extreme digits (8s, 9s -> -7 with carries, 0xf...),
    the generator with z == 1, and the point at infinity (reference ladder's own (x, y, 0) representative)."""
    p, q = _points(5)
    n0 = p.size // G1
    R = o.R_ORDER
    ks = [0, 1, 2, R - 1, 8, 9, 16, 0x88888888888888888888888888888888 % R, int("9" * 60, 16) % R, int("f" * 63, 16) % R,
          int("7" * 63, 16), (1 << 253) + 8, R - 2, R // 2, 0x8000000000000000]
    ks += [rng.randrange(R) for _ in range(17)]
    n = len(ks)
    P = np.tile(p.reshape(n0, G1), (n // n0 + 1, 1))[:n].copy()
    Q = np.tile(q.reshape(n0, G2), (n // n0 + 1, 1))[:n].copy()
    P[5] = np.frombuffer(o.to_bytes((1, 2, 1)), dtype=np.uint8)                                 # generator, z == one
    P[6] = np.frombuffer(o.to_bytes((0, 1, 0)), dtype=np.uint8)                                 # infinity
    P[7] = np.frombuffer(o.to_bytes((5, 7, 0)), dtype=np.uint8)                                 # infinity, odd representative
    Q[6] = np.frombuffer(o.to_bytes(((0, 0), (1, 0), (0, 0))), dtype=np.uint8)
    P, Q = P.reshape(-1), Q.reshape(-1)
    fr = np.concatenate([np.frombuffer(o.fr_to_bytes(k), dtype=np.uint8) for k in ks]).copy()
    # the recoding reconstructs the scalar with digits in -7..8
    for k in ks:
        d = (ctypes.c_byte * 64)()
        hostsim.lib().hs_wnaf4_fixed_recode(k.to_bytes(32, "little"), d)
        assert all(-7 <= x <= 8 for x in d) and sum(int(x) << (4 * i) for i, x in enumerate(d)) == k
    out = np.empty_like(P)
    hostsim.lib().hs_g1_mul_normalized(_p(P), _p(fr), _p(out), ctypes.c_size_t(n))
    assert out.tobytes() == c.g1_normalize(c.g1_mul(P, fr)).tobytes()
    out = np.empty_like(Q)
    hostsim.lib().hs_g2_mul_normalized(_p(Q), _p(fr), _p(out), ctypes.c_size_t(n))
    assert out.tobytes() == c.g2_normalize(c.g2_mul(Q, fr)).tobytes()


def test_precompute_miller_pairing(golden):
    p, q = _points(3)
    n = p.size // G1
    pc = np.zeros(c.G2PRECOMP, dtype=np.uint8)
    hostsim.lib().hs_g2_precompute.restype = ctypes.c_int
    assert hostsim.lib().hs_g2_precompute(_p(q), _p(pc)) == 1
    assert pc.tobytes() == c.g2_precompute(q[:G2]).tobytes()
    out = np.empty(n * FQ12, dtype=np.uint8)
    ok = np.empty(n, dtype=np.uint8)
    hostsim.lib().hs_miller_loop(_p(q), _p(p), _p(out), _p(ok), ctypes.c_size_t(n))
    ro, rok = c.miller_loop_many(q, p)
    assert (ok == rok).all() and out.tobytes() == ro.tobytes()
    hostsim.lib().hs_pairing(_p(p), _p(q), _p(out), ctypes.c_size_t(n))
    assert out.tobytes() == c.pairing_many(p, q).tobytes()
    # the reference's own golden pairing
    v = golden["test_reduced_pairing"]
    g1 = _b(o.g_mul(o.FQ_OPS, o.G1_ONE, int(v["g1_scalar"])))
    g2 = _b(o.g_mul(o.FQ2_OPS, o.G2_ONE, int(v["g2_scalar"])))
    one = np.empty(FQ12, dtype=np.uint8)
    hostsim.lib().hs_pairing(_p(g1), _p(g2), _p(one), ctypes.c_size_t(1))
    e = [int(x) for x in v["expected"]]
    assert one.tobytes() == b"".join(o.fq_to_mont(x).to_bytes(32, "little") for x in e)


def test_gt_pow():
    a = rand_elems(12, 2)
    ks = [rng.randrange(o.R_ORDER), 0]
    fr = np.concatenate([np.frombuffer(o.fr_to_bytes(k), dtype=np.uint8) for k in ks]).copy()
    out = np.empty_like(a)
    hostsim.lib().hs_gt_pow(_p(a), _p(fr), _p(out), ctypes.c_size_t(2))
    assert out.tobytes() == c.gt_pow(a, fr).tobytes()


def test_constants_match_reference_literals(golden):
    m = golden["montgomery_literals"]
    buf = np.zeros(4096, dtype=np.uint8)
    hostsim.lib().hs_constants(_p(buf))
    raw = buf.tobytes()
    vals = [int.from_bytes(raw[32 * i:32 * i + 32], "little") for i in range(40)]
    exp = [int(m["fq"]["one"]), int(m["fq"]["rsquared"]), int(m["fq"]["rcubed"]), int(m["two_inv"]), int(m["g1_coeff_b"]),
           int(m["g1_one_y"])] + [int(x) for x in m["g2_coeff_b"]] + [int(x) for x in m["twist_mul_by_q_x"]] + \
          [int(x) for x in m["twist_mul_by_q_y"]]
    for tab in ("fq6_frobenius_c1", "fq6_frobenius_c2", "fq12_frobenius_c1"):
        for k in ("1", "2", "3"):
            exp += [int(x) for x in m[tab][k]]
    exp += [int(x) for x in m["g2_one"]["x"]] + [int(x) for x in m["g2_one"]["y"]]
    assert vals[:len(exp)] == exp


# ---- bn_wire.cuh against the oracles ----
def test_wire_sqrt_and_validation_host_build():
    from tests import util
    from tests.test_wire_oracle import _mont
    vals = [rng.randrange(o.Q) for _ in range(16)] + [0, 1, 4, o.Q - 1]
    a = _mont(vals)
    out, some = np.empty_like(a), np.empty(len(vals), dtype=np.uint8)
    hostsim.lib().hs_fq_sqrt(_p(a), _p(out), _p(some), ctypes.c_size_t(len(vals)))
    ro, so = c.fq_sqrt(a)
    assert some.tobytes() == so.tobytes()
    for i in range(len(vals)):
        if some[i]:
            assert out[32 * i:32 * i + 32].tobytes() == ro[32 * i:32 * i + 32].tobytes()
    pairs = [(rng.randrange(o.Q), rng.randrange(o.Q)) for _ in range(8)] + [(0, 0), (4, 0), (0, 1), (o.Q - 1, 0), (o.Q - 4, 0), (1, 2)]
    a = _mont([w for pr in pairs for w in pr])
    out, some = np.empty_like(a), np.empty(len(pairs), dtype=np.uint8)
    hostsim.lib().hs_fq2_sqrt(_p(a), _p(out), _p(some), ctypes.c_size_t(len(pairs)))
    ro, so = c.fq2_sqrt(a)
    assert some.tobytes() == so.tobytes()
    for i in range(len(pairs)):
        if some[i]:
            assert out[64 * i:64 * i + 64].tobytes() == ro[64 * i:64 * i + 64].tobytes()
    # the decoder's root finder (norm method, any root): same existence answer, and the root squares back to a
    more = [(rng.randrange(o.Q), rng.randrange(o.Q)) for _ in range(40)] + [(rng.randrange(o.Q), 0) for _ in range(8)] + [(0, rng.randrange(o.Q)) for _ in range(4)]
    a = _mont([w for pr in pairs + more for w in pr])
    npairs = len(pairs) + len(more)
    out, some = np.empty_like(a), np.empty(npairs, dtype=np.uint8)
    hostsim.lib().hs_fq2_sqrt_any(_p(a), _p(out), _p(some), ctypes.c_size_t(npairs))
    ro, so = c.fq2_sqrt(a)
    assert some.tobytes() == so.tobytes() and 0 < int(some.sum()) < npairs
    sq = c.fq2_sqr(out)
    for i in range(npairs):
        if some[i]:
            assert sq[64 * i:64 * i + 64].tobytes() == a[64 * i:64 * i + 64].tobytes()
            root = o.fq_list_from_bytes(ro[64 * i:64 * i + 64].tobytes())
            neg = util.fq_mont_bytes([(-root[0]) % o.Q, (-root[1]) % o.Q])
            assert out[64 * i:64 * i + 64].tobytes() in (ro[64 * i:64 * i + 64].tobytes(), neg)
    g1 = util.affine_bytes([(1, 2), (1, 1), (1, 3), (0, 0)])
    st = np.empty(4, dtype=np.uint8)
    hostsim.lib().hs_g1_affine_check(_p(g1), _p(st), ctypes.c_size_t(4))
    assert list(st) == [0, 1, 1, 1]
    gen = (o.G2_ONE[0], o.G2_ONE[1])
    pts = [gen, ((1, 2), (3, 4))] + util.g2_curve_points_outside_subgroup(2) + util.g2_small_order_points(1)
    pts += [o.g_to_affine(o.FQ2_OPS, o.g_mul(o.FQ2_OPS, o.G2_ONE, k)) for k in (2, o.BN_U, o.R_ORDER - 1, o.BN_U + 1)]
    g2 = util.affine_bytes(pts)
    for ref_ladder in (0, 1):                     # psi test and the reference's [r-1]P + P ladder
        st = np.empty(len(pts), dtype=np.uint8)
        hostsim.lib().hs_g2_affine_check(_p(g2), _p(st), ctypes.c_size_t(len(pts)), ctypes.c_int(ref_ladder))
        assert st.tobytes() == c.g2_affine_check(g2).tobytes() and list(st) == [0, 1, 2, 2, 2, 0, 0, 0, 0]


def test_wire_from_compressed_host_build():
    from tests.test_wire_oracle import wire_cases_g1, wire_cases_g2
    for cases, rec, unit, hs, ref in ((wire_cases_g1(), 33, G1, "hs_g1_from_compressed", c.g1_from_compressed),
                                      (wire_cases_g2(), 65, G2, "hs_g2_from_compressed", c.g2_from_compressed)):
        enc = np.frombuffer(b"".join(e for e, _ in cases), dtype=np.uint8).copy()
        n = len(cases)
        ro, so = ref(enc)
        for extra in ([[]] if unit == G1 else [[ctypes.c_int(0)], [ctypes.c_int(1)]]):
            out, st = np.empty(n * unit, dtype=np.uint8), np.empty(n, dtype=np.uint8)
            getattr(hostsim.lib(), hs)(_p(enc), _p(out), _p(st), ctypes.c_size_t(n), *extra)
            assert st.tobytes() == so.tobytes()
            assert out.tobytes() == ro.tobytes()
            assert set(st) >= {0, 1, 2, 3}


def test_lazy_sums_of_the_quad_lane_executor():
    """acc9_add / acc9_add9 / acc9_reduce (csrc/bn_fq.cuh; the device side ends in fq_red9_ptx, swept by tools/gen_fq_ptx.py): up to
    8 signed terms, optionally plus 9 times a second signed sum, reduced once -- against Python integers, extremes included."""
    lib = hostsim.lib()

    def le(v):
        return np.frombuffer(int(v).to_bytes(32, "little"), dtype=np.uint8)

    def run(ts, signs, xs, xsigns):
        t = np.concatenate([le(v) for v in ts] + [np.zeros(0, np.uint8)]) if ts else np.zeros(32, np.uint8)
        x = np.concatenate([le(v) for v in xs]) if xs else np.zeros(32, np.uint8)
        out = np.zeros(32, np.uint8)
        lib.hs_acc9_sum(_p(t), len(ts), signs, _p(x), len(xs), xsigns, _p(out))
        exp = sum(-v if (signs >> k) & 1 else v for k, v in enumerate(ts)) + 9 * sum(-v if (xsigns >> k) & 1 else v for k, v in enumerate(xs))
        assert int.from_bytes(out.tobytes(), "little") == exp % o.Q, (ts, signs, xs, xsigns)

    big = o.Q - 1
    for n in range(0, 9):
        run([big] * n, 0, [], 0)                    # 8 (q - 1), the largest plain sum
        run([big] * n, (1 << n) - 1, [], 0)         # -8 (q - 1), the smallest
        for nx in range(1, 9 - n):
            run([big] * n, 0, [big] * nx, 0)
            run([big] * n, (1 << n) - 1, [big] * nx, (1 << nx) - 1)
            run([0] * n, 0, [0] * nx, 0)
    for _ in range(3000):
        n = rng.randrange(0, 9)
        nx = rng.randrange(0, 9 - n)
        run([rfq() for _ in range(n)], rng.randrange(1 << n) if n else 0, [rfq() for _ in range(nx)], rng.randrange(1 << nx) if nx else 0)


def test_cyclotomic_out_c_lazy_sum():
    """cyc_outc (csrc/bn_fq.cuh): 3 (S +- m' - 10 m) - 2 z in one lazily accumulated sum, against Python integers (extremes first)."""
    lib = hostsim.lib()
    big = o.Q - 1
    cases = [(big, 0, big, 0), (0, big, 0, big), (big, big, big, big), (0, 0, 0, 0), (big, 0, 0, big), (0, big, big, 0)]
    cases += [(rfq(), rfq(), rfq(), rfq()) for _ in range(3000)]
    n = len(cases)
    arr = [np.frombuffer(b"".join(int(c_[k]).to_bytes(32, "little") for c_ in cases), dtype=np.uint8).copy() for k in range(4)]
    for plus in (0, 1):
        out = np.zeros(32 * n, np.uint8)
        lib.hs_cyc_outc(_p(arr[0]), _p(arr[1]), _p(arr[2]), ctypes.c_int(plus), _p(arr[3]), _p(out), ctypes.c_size_t(n))
        for i, (s_, m_, mx_, z_) in enumerate(cases):
            exp = (3 * (s_ + (mx_ if plus else -mx_) - 10 * m_) - 2 * z_) % o.Q
            assert int.from_bytes(out[32 * i:32 * i + 32].tobytes(), "little") == exp, (i, plus)
