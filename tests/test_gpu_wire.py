"""SURVEY 8f rows N2/N3 on the GPU, through the C ABI: AffineG::new validation and the compressed wire formats
against the C oracle (bit-exact images, identical status bytes), the reference's golden vectors, and a
compress -> decode round trip at a batch size that fills the device."""
import numpy as np
import pytest

from oracle import bn254_ref as o
from oracle import c_oracle as c
from tests import util
from tests.test_wire_oracle import G1_HEX, G1_XY, G2_HEX, G2_X, G2_Y, wire_cases_g1, wire_cases_g2

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def eng():
    import paritytech_bn_b200 as bn
    e = bn.Engine(0)
    yield e
    e.close()


def test_affine_new_reference_cases(eng):
    """affine_fail / affine_ok, src/groups/mod.rs:1138-1154."""
    assert list(eng.g1_affine_check(util.affine_bytes([(1, 2), (1, 1), (1, 3), (0, 0)]))) == [0, 1, 1, 1]
    gen = (o.G2_ONE[0], o.G2_ONE[1])
    assert list(eng.g2_affine_check(util.affine_bytes([gen]))) == [0]
    assert eng.g1_affine_check(np.empty(0, dtype=np.uint8)).size == 0


def test_affine_check_vs_oracle(eng):
    ks = util.scalars(300, stream=41)
    g1, g2 = util.points_cpu(ks, ks[::-1])
    a1 = util.affine_of(c.g1_normalize(g1), 96)
    a2 = util.affine_of(c.g2_normalize(g2), 192)
    # corrupt every third point: flip one bit of y
    a1.reshape(-1, 64)[::3, 40] ^= 1
    a2.reshape(-1, 128)[::3, 70] ^= 4
    off = util.affine_bytes(util.g2_curve_points_outside_subgroup(24, seed=42) + util.g2_small_order_points(8, seed=43))
    a2 = np.concatenate([a2, off])
    st1, st2 = eng.g1_affine_check(a1), eng.g2_affine_check(a2)
    assert st1.tobytes() == c.g1_affine_check(a1).tobytes()
    assert st2.tobytes() == c.g2_affine_check(a2).tobytes()
    assert list(st1[:3]) == [1, 0, 0] and set(st2[-32:]) == {2} and set(st2[:300]) == {0, 1}


def test_from_compressed_golden_vectors(eng):
    """g1_from_compressed / g2_from_compressed, src/lib.rs:681-736."""
    out, st = eng.g1_from_compressed(bytes.fromhex(G1_HEX))
    assert st[0] == 0 and out.tobytes() == util.fq_mont_bytes(list(G1_XY) + [1])
    out, st = eng.g2_from_compressed(bytes.fromhex("0a" + G2_HEX + "0b" + G2_HEX + "0c" + G2_HEX))
    assert list(st) == [0, 0, 1]
    assert out[:192].tobytes() == util.fq_mont_bytes(list(G2_X) + list(G2_Y) + [1, 0])
    assert out[192:384].tobytes() == util.fq_mont_bytes(list(G2_X) + list(o.fq2_neg(G2_Y)) + [1, 0])
    assert not out[384:].any()


def test_from_compressed_every_exit_vs_oracle(eng):
    for cases, fn, ref in ((wire_cases_g1(), eng.g1_from_compressed, c.g1_from_compressed),
                           (wire_cases_g2(), eng.g2_from_compressed, c.g2_from_compressed)):
        enc = b"".join(e for e, _ in cases)
        out, st = fn(enc)
        ro, so = ref(enc)
        assert st.tobytes() == so.tobytes() and out.tobytes() == ro.tobytes()
        assert set(st) == {0, 1, 2, 3}


def _compress_many(eng, n, stream):
    """n (G1, G2) points from the CUDA scalar-multiplication kernels, compressed on the host."""
    ks = util.scalars(n, stream=stream)
    g1 = eng.g1_normalize(eng.g1_mul(np.tile(util.G1_GEN, n), util.fr_bytes(ks)))
    g2 = eng.g2_normalize(eng.g2_mul(np.tile(util.G2_GEN, n), util.fr_bytes(ks[::-1])))
    e1, e2 = [], []
    for i in range(n):
        x, y, _ = o.g1_from_bytes(g1[96 * i:96 * i + 96].tobytes())
        e1.append(util.compress_g1(x, y))
        x, y, _ = o.g2_from_bytes(g2[192 * i:192 * i + 192].tobytes())
        e2.append(util.compress_g2(x, y))
    return g1, g2, b"".join(e1), b"".join(e2)


def test_compress_decode_round_trip(eng):
    """Size-independent property at a device-filling batch: decode(compress(P)) == P, status Ok; a sample is
    also checked against the oracle, and random x coordinates agree with it on which records decode."""
    n = 4096
    g1, g2, e1, e2 = _compress_many(eng, n, 51)
    out1, st1 = eng.g1_from_compressed(e1)
    out2, st2 = eng.g2_from_compressed(e2)
    assert not st1.any() and not st2.any()
    assert out1.tobytes() == g1.tobytes() and out2.tobytes() == g2.tobytes()
    ro, so = c.g2_from_compressed(e2[:65 * 64])
    assert out2[:192 * 64].tobytes() == ro.tobytes() and not so.any()
    rng = np.random.default_rng(52)
    r1 = b"".join(bytes([2 + (i & 1)]) + (util._rand_fq(rng)).to_bytes(32, "big") for i in range(512))
    out, st = eng.g1_from_compressed(r1)
    ro, so = c.g1_from_compressed(r1)
    assert st.tobytes() == so.tobytes() and out.tobytes() == ro.tobytes() and 0 < int((st == 2).sum()) < 512
    r2 = b"".join(bytes([10 + (i & 1)]) + (util._rand_fq(rng) * o.Q + util._rand_fq(rng)).to_bytes(64, "big") for i in range(96))
    out, st = eng.g2_from_compressed(r2)
    ro, so = c.g2_from_compressed(r2)
    assert st.tobytes() == so.tobytes() and out.tobytes() == ro.tobytes() and set(st) == {2}


def test_psi_subgroup_test_matches_reference_ladder(eng):
    """The default 63-bit psi test and the reference's 254-bit [r-1]P + P ladder (BN254_SUBGROUP_REF=1) give the
    same status on members, random curve points, cofactor-order points and off-curve points."""
    import os
    import paritytech_bn_b200 as bn
    ks = util.scalars(500, stream=71) + [1, 2, o.BN_U, o.BN_U + 1, o.R_ORDER - 1]
    g2 = c.g2_normalize(c.g2_mul(np.tile(util.G2_GEN, len(ks)), util.fr_bytes(ks)))
    a2 = util.affine_of(g2, 192)
    a2.reshape(-1, 128)[::7, 100] ^= 8
    neg = util.affine_bytes(util.g2_curve_points_outside_subgroup(200, seed=72) + util.g2_small_order_points(40, seed=73))
    a2 = np.concatenate([a2, neg])
    fast = eng.g2_affine_check(a2)
    os.environ["BN254_SUBGROUP_REF"] = "1"
    try:
        ref_eng = bn.Engine(0)
    finally:
        del os.environ["BN254_SUBGROUP_REF"]
    try:
        slow = ref_eng.g2_affine_check(a2)
        enc = b"".join(e for e, _ in wire_cases_g2())
        o1, s1 = eng.g2_from_compressed(enc)
        o2, s2 = ref_eng.g2_from_compressed(enc)
    finally:
        ref_eng.close()
    assert fast.tobytes() == slow.tobytes() == c.g2_affine_check(a2).tobytes()
    assert set(fast[-240:]) == {2} and int((fast == 0).sum()) > 400
    assert s1.tobytes() == s2.tobytes() and o1.tobytes() == o2.tobytes()


def test_every_entry_point_at_a_ragged_size():
    """scripts/all_entry_points.py: each host-buffer entry point at n = 70 (block size 64) against the oracle."""
    import os
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    p = subprocess.run([sys.executable, os.path.join(root, "scripts", "all_entry_points.py"), "70"], capture_output=True, text=True,
                       timeout=900, cwd=root)
    assert p.returncode == 0 and "every entry point ran" in p.stdout, p.stdout[-1000:] + p.stderr[-3000:]


def test_decode_validate_pair_chain_stays_on_device(eng):
    """Compressed records -> from_compressed_dev -> affine_check_dev -> pairing_many_dev without leaving HBM; the Gt
    results equal the oracle's pairings of the original points, a corrupted record decodes to the all-zero image
    (z = 0), which the pairing treats as the point at infinity (Gt::one), exactly like pairing(G1::zero(), q)."""
    import torch
    n = 200
    g1, g2, e1, e2 = _compress_many(eng, n, 91)
    e2 = bytearray(e2)
    e2[65 * 7] = 0x0c                                           # invalid tag on record 7
    dev = torch.device("cuda", 0)
    stream = torch.cuda.ExternalStream(eng.stream, device=dev)
    d_e1 = torch.frombuffer(bytearray(e1), dtype=torch.uint8).to(dev)
    d_e2 = torch.frombuffer(e2, dtype=torch.uint8).to(dev)
    d_g1 = torch.empty(n * 96, dtype=torch.uint8, device=dev)
    d_g2 = torch.empty(n * 192, dtype=torch.uint8, device=dev)
    d_s1 = torch.empty(n, dtype=torch.uint8, device=dev)
    d_s2 = torch.empty(n, dtype=torch.uint8, device=dev)
    d_gt = torch.empty(n * 384, dtype=torch.uint8, device=dev)
    torch.cuda.synchronize()
    eng.g1_from_compressed_dev(d_e1.data_ptr(), d_g1.data_ptr(), d_s1.data_ptr(), n)
    eng.g2_from_compressed_dev(d_e2.data_ptr(), d_g2.data_ptr(), d_s2.data_ptr(), n)
    eng.pairing_dev(d_g1.data_ptr(), d_g2.data_ptr(), d_gt.data_ptr(), n)
    eng.synchronize()
    s1, s2 = d_s1.cpu().numpy(), d_s2.cpu().numpy()
    assert not s1.any() and s2[7] == 1 and not np.delete(s2, 7).any()
    g2x = g2.copy()
    g2x[192 * 7:192 * 8] = 0
    assert d_g2.cpu().numpy().tobytes() == g2x.tobytes() and d_g1.cpu().numpy().tobytes() == g1.tobytes()
    got = d_gt.cpu().numpy()
    g2inf = g2.copy()
    g2inf[192 * 7:192 * 8] = util.G2_INF
    assert got.tobytes() == c.pairing_many(g1, g2inf).tobytes()
    assert got[384 * 7:384 * 8].tobytes() == util.GT_ONE.tobytes()
    # validation twin on the decoded affine coordinates
    xy = d_g2.view(n, 192)[:, :128].contiguous()
    d_st = torch.empty(n, dtype=torch.uint8, device=dev)
    torch.cuda.synchronize()
    eng.g2_affine_check_dev(xy.data_ptr(), d_st.data_ptr(), n)
    eng.synchronize()
    st = d_st.cpu().numpy()
    assert st[7] == 1 and not np.delete(st, 7).any()         # (0, 0) is not on the curve
