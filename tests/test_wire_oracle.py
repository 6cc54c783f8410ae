"""SURVEY 8f rows N2 (AffineG::new validation) and N3 (sqrt, compressed wire formats): pin the Python and C
oracles to the reference's golden vectors (src/lib.rs:681-736, src/fields/fp.rs:289-296,
src/fields/fq2.rs:235-260, src/groups/mod.rs:1138-1154) and to each other."""
import numpy as np

from oracle import bn254_ref as o
from oracle import c_oracle as c
from tests import util

G2_HEX = "023aed31b5a9e486366ea9988b05dba469c6206e58361d9c065bbea7d928204a761efc6e4fa08ed227650134b52c7f7dd0463963e8a4bf21f4899fe5da7f984a"
G2_X = (5923585509243758863255447226263146374209884951848029582715967108651637186684,
        5336385337059958111259504403491065820971993066694750945459110579338490853570)
G2_Y = (10374495865873200088116930399159835104695426846400310764827677226300185211748,
        5256529835065685814318509161957442385362539991735248614869838648137856366932)
G1_HEX = "0230644e72e131a029b85045b68181585d97816a916871ca8d3c208c16d87cfd46"
G1_XY = (21888242871839275222246405745257275088696311157297823662689037894645226208582,
         3969792565221544645472939191694882283483352126195956956354061729942568608776)


def _mont(vals):
    return np.frombuffer(util.fq_mont_bytes(vals), dtype=np.uint8).copy()


def test_sqrt_golden_vectors():
    """sqrt_fq (src/fields/fp.rs:289-296) and sqrt_fq2 (src/fields/fq2.rs:235-260), both oracles."""
    root = 5204065062716160319596273903996315000119019512886596366359652578430118331601
    assert o.fq_sqrt(348579348568) == root
    out, some = c.fq_sqrt(_mont([348579348568]))
    assert some[0] == 1 and out.tobytes() == _mont([root]).tobytes()
    x1 = (12844195307879678418043983815760255909500142247603239203345049921980497041944,
          7476417578426924565731404322659619974551724117137577781074613937423560117731)
    x2 = (3345897230485723946872934576923485762803457692345760237495682347502347589474,
          1234912378405347958234756902345768290345762348957605678245967234857634857676)
    assert o.fq2_sqrt(x2) == x1
    assert o.fq2_sqrt((o.Q - 1, 0)) == (0, 1)          # i is sqrt(-1)
    assert o.fq2_sqrt((1, 2)) is None                  # no sqrt for 1 + 2i
    out, some = c.fq2_sqrt(_mont(list(x2) + [o.Q - 1, 0, 1, 2]))
    assert list(some) == [1, 1, 0]
    assert out[:128].tobytes() == _mont(list(x1) + [0, 1]).tobytes()


def test_sqrt_c_vs_python_random():
    rng = np.random.default_rng(5)
    vals = [util._rand_fq(rng) for _ in range(24)] + [0, 1, o.Q - 1]
    out, some = c.fq_sqrt(_mont(vals))
    for i, v in enumerate(vals):
        r = o.fq_sqrt(v)
        assert bool(some[i]) == (r is not None)
        if r is not None:
            assert out[32 * i:32 * i + 32].tobytes() == _mont([r]).tobytes()
            assert r * r % o.Q == v
    pairs = [(util._rand_fq(rng), util._rand_fq(rng)) for _ in range(8)] + [(0, 0), (4, 0), (0, 1), (o.Q - 4, 0)]
    out, some = c.fq2_sqrt(_mont([w for p in pairs for w in p]))
    for i, v in enumerate(pairs):
        r = o.fq2_sqrt(v)
        assert bool(some[i]) == (r is not None)
        if r is not None:
            assert out[64 * i:64 * i + 64].tobytes() == _mont(list(r)).tobytes()
            assert o.fq2_sqr(r) == v


def test_affine_new_reference_cases():
    """affine_fail / affine_ok (src/groups/mod.rs:1138-1154): (1, 1) and (1, b) are both rejected; (1, 2) is G1::one."""
    st = c.g1_affine_check(util.affine_bytes([(1, 2), (1, 1), (1, 3), (0, 0)]))
    assert list(st) == [0, 1, 1, 1]
    assert [o.affine_on_curve(o.FQ_OPS, o.G1_B, x, y) for x, y in [(1, 2), (1, 1), (1, 3), (0, 0)]] == [True, False, False, False]
    gen = (o.G2_ONE[0], o.G2_ONE[1])
    assert list(c.g2_affine_check(util.affine_bytes([gen]))) == [0]


def test_g2_subgroup_check_c_vs_python():
    gen = (o.G2_ONE[0], o.G2_ONE[1])
    good = [o.g_to_affine(o.FQ2_OPS, o.g_mul(o.FQ2_OPS, o.G2_ONE, k)) for k in util.scalars(3, stream=21)]
    bad_curve = [((1, 2), (3, 4)), (gen[0], o.fq2_add(gen[1], (1, 0)))]
    off = util.g2_curve_points_outside_subgroup(3)
    small = util.g2_small_order_points(2)
    pts = [gen] + good + bad_curve + off + small
    want = [0] * 4 + [1] * 2 + [2] * (len(off) + len(small))
    got_py = []
    for x, y in pts:
        if not o.affine_on_curve(o.FQ2_OPS, o.G2_B, x, y):
            got_py.append(1)
        else:
            got_py.append(0 if o.g2_in_subgroup(x, y) else 2)
    assert got_py == want
    assert list(c.g2_affine_check(util.affine_bytes(pts))) == want


def test_from_compressed_golden_vectors():
    """g1_from_compressed / g2_from_compressed, src/lib.rs:681-736."""
    st, g = o.g1_from_compressed(bytes.fromhex(G1_HEX))
    assert st == 0 and g == G1_XY + (1,)
    out, stc = c.g1_from_compressed(bytes.fromhex(G1_HEX))
    assert stc[0] == 0 and out.tobytes() == util.fq_mont_bytes(list(G1_XY) + [1])
    st, g = o.g2_from_compressed(bytes.fromhex("0a" + G2_HEX))
    assert st == 0 and g[0] == G2_X and g[1] == G2_Y
    st, g = o.g2_from_compressed(bytes.fromhex("0b" + G2_HEX))            # 0b prefix is the reflected point
    assert st == 0 and g[0] == G2_X and o.fq2_neg(g[1]) == G2_Y
    assert o.g2_from_compressed(bytes.fromhex("0c" + G2_HEX))[0] == o.DECODE_INVALID_ENCODING
    out, stc = c.g2_from_compressed(bytes.fromhex("0a" + G2_HEX + "0b" + G2_HEX + "0c" + G2_HEX))
    assert list(stc) == [0, 0, 1]
    yneg = o.fq2_neg(G2_Y)
    assert out[:192].tobytes() == util.fq_mont_bytes(list(G2_X) + list(G2_Y) + [1, 0])
    assert out[192:384].tobytes() == util.fq_mont_bytes(list(G2_X) + list(yneg) + [1, 0])
    assert not out[384:].any()


def wire_cases_g1():
    """(encoding, expected status) pairs covering every exit of G1::from_compressed."""
    cases = []
    for k in util.scalars(6, stream=31):
        x, y, _ = o.g_normalize(o.FQ_OPS, o.g_mul(o.FQ_OPS, o.G1_ONE, k))
        cases.append((util.compress_g1(x, y), 0))
        cases.append((util.compress_g1(x, (-y) % o.Q), 0))
    cases.append((bytes([2]) + (o.Q).to_bytes(32, "big"), 3))              # x = q: Field(NotMember)
    cases.append((bytes([7]) + (2**256 - 1).to_bytes(32, "big"), 3))        # field check precedes the tag check
    x = 0
    while o.fq_sqrt((x**3 + 3) % o.Q) is not None:
        x += 1
    cases.append((bytes([2]) + x.to_bytes(32, "big"), 2))                  # no root: NotMember
    cases.append((bytes([9]) + x.to_bytes(32, "big"), 2))                  # ... even with a bad tag (root first)
    cases.append((bytes([4]) + (1).to_bytes(32, "big"), 1))                # x = 1 has the root 2: bad tag
    cases.append((bytes([0]) + (1).to_bytes(32, "big"), 1))
    return cases


def wire_cases_g2():
    cases = []
    for k in util.scalars(3, stream=32):
        x, y = o.g_to_affine(o.FQ2_OPS, o.g_mul(o.FQ2_OPS, o.G2_ONE, k))
        cases.append((util.compress_g2(x, y), 0))
        cases.append((util.compress_g2(x, o.fq2_neg(y)), 0))
    (x, y), = util.g2_curve_points_outside_subgroup(1, seed=33)
    cases.append((util.compress_g2(x, y), 2))                              # on the curve, outside the subgroup
    cases.append((bytes([10]) + (o.Q * o.Q).to_bytes(64, "big"), 3))       # quotient = q: Field(NotMember)
    cases.append((bytes([10]) + (2**512 - 1).to_bytes(64, "big"), 3))      # quotient >= 2^256
    cases.append((bytes([10]) + (o.Q * o.Q - 1).to_bytes(64, "big"), None))  # largest valid x = (q-1, q-1)
    xx = (0, 0)
    while o.fq2_sqrt(o.fq2_add(o.fq2_mul(o.fq2_sqr(xx), xx), o.G2_B)) is not None:
        xx = (xx[0] + 1, 0)
    cases.append((bytes([10]) + (xx[1] * o.Q + xx[0]).to_bytes(64, "big"), 2))
    cases.append((bytes([12]) + (xx[1] * o.Q + xx[0]).to_bytes(64, "big"), 2))  # root is tested before the tag
    gx = o.G2_ONE[0]
    cases.append((bytes([0]) + (gx[1] * o.Q + gx[0]).to_bytes(64, "big"), 1))
    return cases


def test_from_compressed_c_vs_python_all_exits():
    cases = wire_cases_g1()
    out, st = c.g1_from_compressed(b"".join(e for e, _ in cases))
    for i, (enc, want) in enumerate(cases):
        s, g = o.g1_from_compressed(enc)
        assert s == st[i] and (want is None or s == want), (i, s, st[i], want)
        img = out[96 * i:96 * i + 96].tobytes()
        assert img == (util.fq_mont_bytes(list(g)) if s == 0 else bytes(96))
    cases = wire_cases_g2()
    out, st = c.g2_from_compressed(b"".join(e for e, _ in cases))
    for i, (enc, want) in enumerate(cases):
        s, g = o.g2_from_compressed(enc)
        assert s == st[i] and (want is None or s == want), (i, s, st[i], want)
        img = out[192 * i:192 * i + 192].tobytes()
        assert img == (util.fq_mont_bytes(list(g[0]) + list(g[1]) + [1, 0]) if s == 0 else bytes(192))


def test_fast_subgroup_test_is_exact():
    """The device's default G2 order check (bn_wire.cuh g2_in_subgroup_psi) replaces [r-1]P + P == 0 by
    [u+1]P + psi([u]P) + psi^2([u]P) == psi^3([2u]P).  Exactness on all of E'(Fq2): psi satisfies
    g(X) = X^2 - tX + q there, so a point passing the test is killed by Res(f, g) with
    f(X) = -2uX^3 + uX^2 + uX + u + 1.  r divides that integer (the test holds on the subgroup) and it is
    coprime to the cofactor 2q - r (nothing else passes)."""
    from math import gcd
    u, q, r = o.BN_U, o.Q, o.R_ORDER
    t = 6 * u * u + 1
    assert q + 1 - t == r
    cof = 2 * q - r                               # #E'(Fq2) = r * (2q - r)
    assert gcd(r, cof) == 1
    # f mod g with X^2 = tX - q, X^3 = (t^2 - q)X - tq; Res(f, g) = c1^2 q + c0 c1 t + c0^2 for monic g
    c1 = -2 * u * (t * t - q) + u * t + u
    c0 = 2 * u * t * q - u * q + (u + 1)
    res = c1 * c1 * q + c0 * c1 * t + c0 * c0
    assert res % r == 0 and gcd(res, cof) == 1
    lam = q % r                                   # psi acts on the subgroup as multiplication by q
    assert (u + 1 + u * lam + u * lam * lam - 2 * u * pow(lam, 3, r)) % r == 0


def _u512(limbs64):
    return sum(v << (64 * i) for i, v in enumerate(limbs64))


def test_u512_divrem_reference_vectors():
    """testing_divrem, src/arith.rs:588-785 (the cases taken modulo the Fq modulus), on the C oracle and on the
    host build of the device routine; then random values against Python's divmod."""
    import ctypes
    from tests import hostsim
    Q = o.Q

    def dev(value):
        be = np.frombuffer(int(value).to_bytes(64, "big"), dtype=np.uint8).copy()
        quot, rem = np.zeros(32, dtype=np.uint8), np.zeros(32, dtype=np.uint8)
        ok = hostsim.lib().hs_u512_divrem_q(be.ctypes.data_as(ctypes.c_void_p), quot.ctypes.data_as(ctypes.c_void_p),
                                            rem.ctypes.data_as(ctypes.c_void_p))
        return (int.from_bytes(quot.tobytes(), "little") if ok else None), int.from_bytes(rem.tobytes(), "little")

    q2m1 = _u512([0x3b5458a2275d69b0, 0xa602072d09eac101, 0x4a50189c6d96cadc, 0x04689e957a1242c8,
                  0x26edfa5c34c6b38d, 0xb00b855116375606, 0x599a6f7c0348d21c, 0x0925c4b8763cbf9c])
    assert q2m1 == Q * Q - 1
    cases = [
        (Q, (1, 0)),                                    # modulus -> 1*q + 0
        (q2m1, (Q - 1, Q - 1)),                         # q^2 - 1 -> (q-1) q + (q-1)
        (q2m1 - 1, (Q - 1, Q - 2)),
        ((1 << 512) - 1, (None, _u512([0xf32cfc5b538afa88, 0xb5e71911d44501fb, 0x47ab1eff0a417ff6, 0x06d89f71cab8351f]))),
        (q2m1 + 1, (None, 0)),                          # q^2 fails
        (q2m1 + 2, (None, 1)),
    ]
    for value, want in cases:
        assert c.u512_divrem_q(value) == want, hex(value)
        assert dev(value) == want, hex(value)
    rng = np.random.default_rng(9)
    for _ in range(100):                                # the reference's loop: c1*q + c0 round trip
        c0, c1 = util._rand_fq(rng), util._rand_fq(rng)
        assert c.u512_divrem_q(c1 * Q + c0) == (c1, c0) and dev(c1 * Q + c0) == (c1, c0)
    for _ in range(100):
        v = int.from_bytes(rng.bytes(64), "little") >> int(rng.integers(0, 260))
        qq, rr = divmod(v, Q)
        want = (qq if qq < Q else None, rr)
        assert c.u512_divrem_q(v) == want and dev(v) == want
