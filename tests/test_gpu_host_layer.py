"""The host layer's promises (include/bn254_b200.h, top comment), exercised on a GPU through the C ABI: concurrent callers of
ONE context serialise safely and callers of distinct contexts run side by side (ctypes releases the GIL during a call); a
pinned output buffer whose registration covers only part of the result falls back to the copy path instead of faulting;
misaligned device pointers, foreign precomp handles and bad arguments are refused with BN254_ERR_ARG; steady-state calls do
not grow the arenas (no cudaMalloc: device memory in use stays flat)."""
import ctypes
import threading

import numpy as np
import pytest

from oracle import c_oracle as c
from tests import util

pytestmark = pytest.mark.gpu

G1, G2, GT = 96, 192, 384


@pytest.fixture(scope="module")
def eng():
    import paritytech_bn_b200 as bn
    e = bn.Engine(0)
    yield e
    e.close()


def _pairs(n, stream):
    return util.points_cpu(util.scalars(n, stream=stream), util.scalars(n, stream=stream + 1))


def test_concurrent_callers_one_context_and_two(eng):
    import paritytech_bn_b200 as bn
    e2 = bn.Engine(0)
    try:
        work = []
        for k, n in enumerate((40, 3000, 7, 2500, 90, 2400)):       # both dispatch paths (warp-cooperative <= 2368 < thread)
            g1, g2 = _pairs(min(n, 128), 801 + 2 * k)
            reps = -(-n // min(n, 128))
            g1 = np.tile(g1, reps)[:n * G1].copy(); g2 = np.tile(g2, reps)[:n * G2].copy()
            work.append((g1, g2, c.pairing_many(g1[:min(n, 128) * G1], g2[:min(n, 128) * G2])))
        results, errors = {}, []

        def run(tid, engine, items):
            try:
                for j, (g1, g2, _) in enumerate(items):
                    for rep in range(3):
                        results[(tid, j, rep)] = engine.pairing(g1, g2)
            except Exception as ex:      # noqa: BLE001
                errors.append(ex)
        # four threads hammer ONE context, two more use a second context at the same time
        ts = [threading.Thread(target=run, args=(t, eng, work[t % 3::3] + work)) for t in range(4)]
        ts += [threading.Thread(target=run, args=(4 + t, e2, work[::-1])) for t in range(2)]
        for t in ts:
            t.start()
        for t in ts:
            t.join()
        assert not errors, errors
        for t in range(6):
            items = (work[t % 3::3] + work) if t < 4 else work[::-1]
            for j, (g1, g2, exp) in enumerate(items):
                m = min(len(g1) // G1, 128)
                for rep in range(3):
                    got = results[(t, j, rep)]
                    assert got[:m * GT].tobytes() == exp.tobytes(), (t, j, rep)
                    assert got.tobytes() == results[(t, j, 0)].tobytes()
    finally:
        e2.close()


def test_partially_pinned_output_falls_back_to_the_copy_path(eng):
    """Zero-copy result stores are taken only when the WHOLE output extent is mapped: register the first half of a
    pageable buffer, then the whole of it -- both calls must give the oracle's bytes and nothing may fault."""
    import torch
    n = 70000                                                       # thread-per-pairing path (zero-copy epilogue lives there)
    base = 256
    p1, p2 = _pairs(base, 821)
    idx = np.arange(n) % base
    g1 = p1.reshape(base, G1)[idx].reshape(-1).copy(); g2 = p2.reshape(base, G2)[(idx * 3 + 1) % base].reshape(-1).copy()
    exp = c.pairing_many(g1[:base * G1], g2[:base * G2])
    out = np.zeros(n * GT, dtype=np.uint8)
    rt = torch.cuda.cudart()
    half = (out.nbytes // 2) & ~4095
    for nbytes in (half, out.nbytes):
        rc = rt.cudaHostRegister(out.ctypes.data, nbytes, 0)
        assert int(rc) == 0, rc
        try:
            out[:] = 0
            got = eng.pairing(g1, g2, out=out)
            assert got[:base * GT].tobytes() == exp.tobytes()
            assert got.reshape(n, GT)[n - 1].tobytes() == exp.reshape(base, GT)[(n - 1) % base].tobytes()   # the unmapped half too
        finally:
            rt.cudaHostUnregister(out.ctypes.data)


def test_argument_checks(eng):
    import torch

    import paritytech_bn_b200 as bn
    dev = torch.device("cuda", 0)
    lib = eng.lib
    g1, g2 = _pairs(4, 831)
    d1 = torch.from_numpy(np.concatenate([np.zeros(16, np.uint8), g1])).to(dev)
    d2 = torch.from_numpy(g2).to(dev)
    do = torch.empty(4 * GT + 16, dtype=torch.uint8, device=dev)
    ok = lib.bn254_pairing_many_dev(eng.ctx, ctypes.c_void_p(d1.data_ptr() + 16), ctypes.c_void_p(d2.data_ptr()), ctypes.c_void_p(do.data_ptr()), ctypes.c_size_t(4))
    assert ok == 0
    eng.synchronize()
    assert do[:4 * GT].cpu().numpy().tobytes() == c.pairing_many(g1, g2).tobytes()
    # 16-byte alignment of device pointers is checked, not assumed
    rc = lib.bn254_pairing_many_dev(eng.ctx, ctypes.c_void_p(d1.data_ptr() + 8), ctypes.c_void_p(d2.data_ptr()), ctypes.c_void_p(do.data_ptr()), ctypes.c_size_t(1))
    assert rc == -1 and b"aligned" in lib.bn254_last_error()
    rc = lib.bn254_pairing_many_dev(eng.ctx, ctypes.c_void_p(d1.data_ptr() + 16), ctypes.c_void_p(d2.data_ptr()), ctypes.c_void_p(do.data_ptr() + 4), ctypes.c_size_t(1))
    assert rc == -1
    # null pointers / contexts, the shared-table limits of the product entry points
    assert lib.bn254_pairing_many(eng.ctx, None, g2.ctypes.data_as(ctypes.c_void_p), do.data_ptr(), ctypes.c_size_t(1)) == -1
    assert lib.bn254_pairing_many(None, g1.ctypes.data_as(ctypes.c_void_p), g2.ctypes.data_as(ctypes.c_void_p), ctypes.c_void_p(0), ctypes.c_size_t(1)) == -1
    g2nine = np.tile(g2[:G2], 9)
    pc9 = eng.g2_precompute(g2nine)
    with pytest.raises(bn.BnError, match="at most 8 shared tables"):
        eng.pairing_batch_precomp(pc9, np.tile(g1[:G1], 9), g1[:G1], g2[:G2], 1)
    pc9.close()
    # a context list naming the same context twice is refused (it would deadlock on its own lock otherwise)
    with pytest.raises(bn.BnError, match="twice"):
        bn.pairing_product_sharded([eng, eng], g1, g2)
    # unknown option names never reach the C struct
    with pytest.raises(KeyError):
        bn.Engine(0, no_such_option=1)


def test_steady_state_calls_do_not_allocate(eng):
    """After one warm-up call per entry point the arenas have their size: repeating the same calls leaves the device's free
    memory untouched (no cudaMalloc / cudaFree on the steady-state path of ANY host entry point)."""
    import torch
    g1, g2 = _pairs(64, 841)
    fr = util.fr_bytes(util.scalars(64, stream=843))
    w = util.groth16_checks(16)
    big1, big2 = np.tile(g1, 60), np.tile(g2, 60)                   # 3840 pairs: the thread-per-pairing kernels and their workspace

    def everything():
        e = eng.pairing(g1, g2); eng.pairing(big1, big2)
        ml, _ = eng.miller_loop_many(g2, g1); eng.final_exponentiation(ml); eng.miller_loop_batch(g2, g1); eng.miller_product(g1, g2)
        eng.gt_mul(e, e); eng.gt_pow(e, fr); eng.gt_inverse(e); eng.gt_eq(e, e)
        eng.g1_mul(g1, fr); eng.g2_mul(g2, fr); eng.g1_mul_normalized(g1, fr); eng.g2_mul_normalized(g2, fr)
        eng.g1_normalize(g1); eng.g2_normalize(g2); eng.g1_add(g1, g1); eng.g2_eq(g2, g2)
        eng.pairing_batch(w["g1"], w["g2"], pairs_per_check=4); eng.pairing_batch(big1, big2, pairs_per_check=12)
        eng.g1_affine_check(util.affine_of(eng.g1_normalize(g1), G1)); eng.g2_affine_check(util.affine_of(eng.g2_normalize(g2), G2))
        eng.field_op("fq12_mul", e, e)

    everything()
    torch.cuda.synchronize()
    free0, _ = torch.cuda.mem_get_info(0)
    for _ in range(3):
        everything()
    torch.cuda.synchronize()
    free1, _ = torch.cuda.mem_get_info(0)
    assert free1 == free0, (free0, free1)


def test_pipeline_lane_counts_give_identical_bytes():
    """bn254_pairing_many splits a host batch of >= 2^16 pairs into 2 .. 8 part batches on as many streams (option
    `pipeline`); every split, pageable or pinned output, must return the bytes of the single-batch path.  Ragged n: the last
    part is not a whole number of blocks."""
    import torch

    import paritytech_bn_b200 as bn
    n, base = (1 << 16) + 700, 256           # one wave (148 x 448 = 66,304): every requested part count is honoured
    p1, p2 = _pairs(base, 861)
    idx = np.arange(n) % base
    g1 = p1.reshape(base, G1)[idx].reshape(-1).copy(); g2 = p2.reshape(base, G2)[(idx * 5 + 2) % base].reshape(-1).copy()
    exp = c.pairing_many(g1[:base * G1], g2[:base * G2])
    ref = None
    pinned = torch.empty(n * GT, dtype=torch.uint8, pin_memory=True).numpy()
    for lanes in (0, 2, 3, 4, 7, 8):
        e = bn.Engine(0, pipeline=lanes)
        try:
            got = e.pairing(g1, g2)
            if ref is None:
                ref = got.copy()
                assert ref[:base * GT].tobytes() == exp.tobytes()
            assert got.tobytes() == ref.tobytes(), lanes
            pinned[:] = 0
            e.pairing(g1, g2, out=pinned)
            assert pinned.tobytes() == ref.tobytes(), lanes
        finally:
            e.close()
