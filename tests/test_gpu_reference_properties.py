"""The reference's own property tests, run on the CUDA path through the C ABI: `test_bilinearity`,
`test_batch_bilinearity_{empty,one,fifty}` (src/groups/mod.rs:1034-1124) and `group_trials` with its four random tests
(src/groups/tests.rs:4-106, instantiated for G1 and G2 at src/groups/mod.rs:481-489).  Same structure and counts (50 trials); the
random stream is this repo's SplitMix64 (StdRng of the `rand` crate is not available here, and the properties hold for any
points).  No oracle is involved: these are algebraic identities between outputs of the engine itself, as in the reference."""
import numpy as np
import pytest

from tests import util

pytestmark = pytest.mark.gpu

G1, G2, GT, FR = 96, 192, 384, 32
R = util.R
N = 50


@pytest.fixture(scope="module")
def eng():
    import paritytech_bn_b200 as bn
    e = bn.Engine(0)
    yield e
    e.close()


def _g1(eng, stream, n=N):
    return eng.g1_mul(np.tile(util.G1_GEN, n), util.fr_bytes(util.scalars(n, stream=stream)))


def _g2(eng, stream, n=N):
    return eng.g2_mul(np.tile(util.G2_GEN, n), util.fr_bytes(util.scalars(n, stream=stream)))


def test_bilinearity(eng):
    """src/groups/mod.rs:1094-1124: a = e(P,Q)^s == e(sP,Q) == e(P,sQ); a != 1; a^(-1 mod r) * a == 1."""
    p, q = _g1(eng, 701), _g2(eng, 702)
    s = util.fr_bytes(util.scalars(N, stream=703))
    sp, sq = eng.g1_mul(p, s), eng.g2_mul(q, s)
    a = eng.gt_pow(eng.pairing(p, q), s)
    b = eng.pairing(sp, q)
    c = eng.pairing(p, sq)
    assert a.tobytes() == b.tobytes() == c.tobytes()
    assert eng.gt_eq(a, b).all() and eng.gt_eq(b, c).all()
    one = np.tile(util.GT_ONE, N)
    assert not eng.gt_eq(a, one).any()
    t = util.fr_bytes([R - 1] * N)                          # -Fr::one()
    assert eng.gt_mul(eng.gt_pow(a, t), a).tobytes() == one.tobytes()


def test_batch_bilinearity_empty(eng):
    """:1034-1040: pairing_batch of nothing is one."""
    e = np.zeros(0, dtype=np.uint8)
    out, is_one = eng.pairing_batch(e, e)
    assert out.tobytes() == util.GT_ONE.tobytes() and is_one[0] == 1


def test_batch_bilinearity_one_and_fifty(eng):
    """:1042-1092: pairing_batch([s P_i], [Q_i]) == pairing_batch([P_i], [s Q_i]) for 1 and for 50 pairs."""
    p, q = _g1(eng, 711), _g2(eng, 712)
    s = util.fr_bytes(util.scalars(N, stream=713))
    sp, sq = eng.g1_mul(p, s), eng.g2_mul(q, s)
    b1, _ = eng.pairing_batch(sp[:G1], q[:G2])
    c1, _ = eng.pairing_batch(p[:G1], sq[:G2])
    assert b1.tobytes() == c1.tobytes() != util.GT_ONE.tobytes()
    b50, _ = eng.pairing_batch(sp, q)
    c50, _ = eng.pairing_batch(p, sq)
    assert b50.tobytes() == c50.tobytes() != util.GT_ONE.tobytes()
    # and the batch equals the product of the single pairings (src/groups/mod.rs:609-640)
    singles = eng.pairing(sp, q).reshape(N, GT)
    acc = singles[0].copy()
    for i in range(1, N):
        acc = eng.gt_mul(acc, singles[i])
    assert acc.tobytes() == b50.tobytes()


@pytest.mark.parametrize("which", ["g1", "g2"])
def test_group_trials(eng, which):
    """src/groups/tests.rs:85-106 and the four random tests it calls (:4-83), 50 trials each, elementwise over the batch."""
    unit, gen, inf = (G1, util.G1_GEN, util.G1_INF) if which == "g1" else (G2, util.G2_GEN, util.G2_INF)
    mul, add, neg, dbl, eq = (getattr(eng, "%s_%s" % (which, op)) for op in ("mul", "add", "neg", "double", "eq"))
    rand = (lambda st: _g1(eng, st)) if which == "g1" else (lambda st: _g2(eng, st))

    def sub(a, b):
        return add(a, neg(b))

    def is_zero(a):
        return (np.asarray(a).reshape(-1, unit)[:, 2 * unit // 3:] == 0).all(axis=1)          # z == 0

    def fr(ks):
        return util.fr_bytes(ks)

    # group_trials preamble (:86-91)
    assert is_zero(inf).all()
    assert is_zero(sub(gen, gen)).all()
    assert eq(add(gen, gen), mul(gen, fr([2]))).all()
    assert is_zero(dbl(inf)).all()
    assert is_zero(add(mul(gen, fr([R - 1])), gen)).all()
    # random_test_addition (:4-16): (a + b) + c == a + (b + c)
    a, b, c = rand(721), rand(722), rand(723)
    assert eq(add(add(a, b), c), add(a, add(b, c))).all()
    # random_test_doubling (:18-28): (a + b) + a == a.double() + b
    assert eq(add(add(a, b), a), add(dbl(a), b)).all()
    # random_test_dh (:30-40): (one * alice_sk) * bob_sk == (one * bob_sk) * alice_sk
    ka, kb = fr(util.scalars(N, stream=724)), fr(util.scalars(N, stream=725))
    one = np.tile(gen, N)
    assert eq(mul(mul(one, ka), kb), mul(mul(one, kb), ka)).all()
    # random_test_equality (:42-83): scramble with a, b, c, d and undo with the inverse scalars
    begin = rand(726)
    ks_a, ks_c = util.scalars(N, stream=727), util.scalars(N, stream=728)
    fa, fc = fr(ks_a), fr(ks_c)
    fai, fci = fr([pow(k, -1, R) for k in ks_a]), fr([pow(k, -1, R) for k in ks_c])
    fti = fr([pow(2, -1, R)] * N)
    bb, dd = rand(729), rand(730)
    acc = begin
    for _ in range(10):
        acc = mul(acc, fa); acc = neg(acc); acc = add(acc, bb); acc = mul(acc, fc); acc = neg(acc); acc = sub(acc, dd); acc = dbl(acc)
    for _ in range(10):
        acc = mul(acc, fti); acc = add(acc, dd); acc = neg(acc); acc = mul(acc, fci); acc = sub(acc, bb); acc = neg(acc); acc = mul(acc, fai)
    assert eq(acc, begin).all()
    # affine <-> Jacobian round trip (src/groups/mod.rs:491-513): normalising twice changes nothing, and == sees the same point
    norm = getattr(eng, which + "_normalize")
    n1 = norm(begin)
    assert norm(n1).tobytes() == n1.tobytes() and eq(n1, begin).all()
