"""G * Fr rates: the reference-sequence kernels (block-compacted additions vs one thread per scalar) and the window kernel,
device-resident, CUDA events on the library's stream.  Output: profiles/r02_bench_scalar_mul.txt"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402
import torch  # noqa: E402

import paritytech_bn_b200 as bn  # noqa: E402
from tests import util  # noqa: E402

dev = torch.device("cuda", 0)
engines = {"compact": bn.Engine(0), "thread": bn.Engine(0, mul_compact=0)}
for which, k, unit in (("g1", 20, 96), ("g2", 18, 192)):
    n = 1 << k
    e = engines["compact"]
    gen = torch.from_numpy(np.tile(util.G1_GEN if which == "g1" else util.G2_GEN, n)).to(dev)
    fr0 = torch.from_numpy(util.fr_images(n, stream=1)).to(dev)
    fr = torch.from_numpy(util.fr_images(n, stream=2)).to(dev)
    base = torch.empty(n * unit, dtype=torch.uint8, device=dev)
    outs = {}
    torch.cuda.synchronize()
    getattr(e, which + "_mul_dev")(gen.data_ptr(), fr0.data_ptr(), base.data_ptr(), n)
    e.synchronize()
    for name, eng in engines.items():
        out = torch.empty(n * unit, dtype=torch.uint8, device=dev)
        stream = torch.cuda.ExternalStream(eng.stream, device=dev)
        fn = getattr(eng, which + "_mul_dev")
        for _ in range(2):
            fn(base.data_ptr(), fr.data_ptr(), out.data_ptr(), n)
        eng.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(stream)
        for _ in range(3):
            fn(base.data_ptr(), fr.data_ptr(), out.data_ptr(), n)
        b.record(stream)
        eng.synchronize()
        ms = a.elapsed_time(b) / 3
        outs[name] = out
        print("%s 2^%d %-8s %8.3f ms  %7.3f M/s" % (which, k, name, ms, n / ms / 1e3), flush=True)
    assert torch.equal(outs["compact"], outs["thread"])
    out = torch.empty(n * unit, dtype=torch.uint8, device=dev)
    stream = torch.cuda.ExternalStream(e.stream, device=dev)
    fn = getattr(e, which + "_mul_normalized_dev")
    fn(base.data_ptr(), fr.data_ptr(), out.data_ptr(), n); e.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record(stream)
    for _ in range(3):
        fn(base.data_ptr(), fr.data_ptr(), out.data_ptr(), n)
    b.record(stream); e.synchronize()
    ms = a.elapsed_time(b) / 3
    print("%s 2^%d %-8s %8.3f ms  %7.3f M/s" % (which, k, "window", ms, n / ms / 1e3), flush=True)
