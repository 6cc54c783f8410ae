"""bn254_final_exp_many_dev_noalloc (no counterpart in the reference): the final exponentiation with its
scratch in a static device array instead of the context's cudaMalloc'ed workspace.  Called through the C ABI and compared
BIT-EXACTLY (integer work: the tolerance of every test here is zero) with the existing entry points, on two calls in a row
that share the static scratch.  Collected last on purpose."""
import ctypes

import numpy as np
import pytest

from tests import util

pytestmark = pytest.mark.gpu

G1, G2, GT = 96, 192, 384
MAX_N = 4096      # BN254_NOALLOC_MAX_N in include/bn254_b200.h


def _noalloc(eng, d_in, d_out, d_some, n):
    return eng.lib.bn254_final_exp_many_dev_noalloc(eng.ctx, ctypes.c_void_p(d_in), ctypes.c_void_p(d_out), ctypes.c_void_p(d_some),
                                                    ctypes.c_size_t(n))


def test_noalloc_final_exponentiation_matches_existing_entry_point_twice():
    import torch

    import paritytech_bn_b200 as bn
    eng = bn.Engine(0)
    try:
        dev = torch.device("cuda", 0)
        # Miller values of random pairs = realistic inputs of the final exponentiation; one zero (Option::None) among them
        n1, n2 = 1500, 333
        ks1, ks2 = util.scalars(n1, stream=601), util.scalars(n1, stream=602)
        g1 = eng.g1_mul(np.tile(util.G1_GEN, n1), util.fr_bytes(ks1))
        g2 = eng.g2_mul(np.tile(util.G2_GEN, n1), util.fr_bytes(ks2))
        ml, ok = eng.miller_loop_many(g2, g1)
        assert ok.all()
        ml = ml.reshape(n1, GT).copy()
        ml[7] = 0
        first = ml.reshape(-1)
        second = ml[::-1][:n2].copy().reshape(-1)          # other values, fewer of them: the scratch still holds call 1's tail
        for call, (vals, n) in enumerate(((first, n1), (second, n2))):
            exp, exp_some = eng.final_exponentiation(vals)                         # existing host entry point
            d_in = torch.from_numpy(vals).to(dev)
            d_out = torch.empty(n * GT, dtype=torch.uint8, device=dev)
            d_some = torch.empty(n, dtype=torch.uint8, device=dev)
            d_ref = torch.empty(n * GT, dtype=torch.uint8, device=dev)
            d_ref_some = torch.empty(n, dtype=torch.uint8, device=dev)
            torch.cuda.synchronize()
            rc = _noalloc(eng, d_in.data_ptr(), d_out.data_ptr(), d_some.data_ptr(), n)
            assert rc == 0, eng.lib.bn254_last_error().decode()
            eng._check(eng.lib.bn254_final_exp_many_dev(eng.ctx, ctypes.c_void_p(d_in.data_ptr()), ctypes.c_void_p(d_ref.data_ptr()),
                                                        ctypes.c_void_p(d_ref_some.data_ptr()), ctypes.c_size_t(n)))   # existing device entry point
            eng.synchronize()
            got, some = d_out.cpu().numpy(), d_some.cpu().numpy()
            assert got.tobytes() == d_ref.cpu().numpy().tobytes() == exp.tobytes(), "call %d" % call
            assert (some == exp_some).all() and (some == d_ref_some.cpu().numpy()).all(), "call %d" % call
            if call == 0:
                assert some[7] == 0 and not got.reshape(n, GT)[7].any()               # None: flag 0, zero bytes
        assert exp_some.sum() == n2                                               # the zero value was in call 1 only
        # in place (d_gt_out == d_gt_in) and without the Option flags
        d_io = torch.from_numpy(second).to(dev)
        torch.cuda.synchronize()
        assert _noalloc(eng, d_io.data_ptr(), d_io.data_ptr(), 0, n2) == 0
        eng.synchronize()
        assert d_io.cpu().numpy().tobytes() == exp.tobytes()
        # the stated limit: one more than the maximum is refused with a clear error, the maximum itself is accepted
        big = torch.from_numpy(np.tile(first[:GT * 64], MAX_N // 64 + 1)).to(dev)
        outb = torch.empty_like(big)
        torch.cuda.synchronize()
        assert _noalloc(eng, big.data_ptr(), outb.data_ptr(), 0, MAX_N + 1) == -1
        assert b"BN254_NOALLOC_MAX_N" in eng.lib.bn254_last_error()
        assert _noalloc(eng, big.data_ptr(), outb.data_ptr(), 0, MAX_N) == 0
        eng.synchronize()
        e64, _ = eng.final_exponentiation(first[:GT * 64])
        assert outb.cpu().numpy()[:MAX_N * GT].tobytes() == np.tile(e64, MAX_N // 64).tobytes()
    finally:
        eng.close()
