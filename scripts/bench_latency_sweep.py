"""Latency of bn254_pairing_many (host buffers, wall clock around the call) against the batch size on the two implementations:
four warps per pairing (k_coop4_pairing, forced with coop4_max_n = 2^20), one warp per pairing (k_coop_pairing, forced with
coop_max_n = 2^20, coop4_max_n = 0) and one thread per pairing (coop_max_n = 0).  Decides the dispatch thresholds coop4_max_n
and coop_max_n.  Output: profiles/r02_latency_sweep.txt"""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402

import paritytech_bn_b200 as bn  # noqa: E402
from tests import util  # noqa: E402

base = 1024
thread = bn.Engine(0, coop_max_n=0)
warp = bn.Engine(0, coop_max_n=1 << 20, coop4_max_n=0)
quad = bn.Engine(0, coop_max_n=1 << 20, coop4_max_n=1 << 20)
p1 = thread.g1_mul(np.tile(util.G1_GEN, base), util.fr_images(base, stream=1)).reshape(base, 96)
p2 = thread.g2_mul(np.tile(util.G2_GEN, base), util.fr_images(base, stream=2)).reshape(base, 192)
print("%8s %12s %12s %12s %12s %12s %12s" % ("n", "quad ms", "warp ms", "thread ms", "quad M/s", "warp M/s", "thread M/s"))
for n in (1, 2, 8, 32, 148, 296, 444, 592, 888, 1184, 2368, 3552, 4736, 7104, 9472, 16384, 32768):
    idx = np.arange(n) % base
    g1, g2 = p1[idx].reshape(-1).copy(), p2[(idx * 7 + 3) % base].reshape(-1).copy()
    res = []
    outs = []
    for e in (quad, warp, thread):
        if e is quad and n > 4736:
            res.append(float('nan')); outs.append(None); continue
        out = e.pairing(g1, g2)
        outs.append(out)
        best = 1e9
        for _ in range(5):
            t0 = time.perf_counter()
            e.pairing(g1, g2)
            best = min(best, time.perf_counter() - t0)
        res.append(best)
    assert outs[1].tobytes() == outs[2].tobytes() and (outs[0] is None or outs[0].tobytes() == outs[2].tobytes())
    print("%8d %12.3f %12.3f %12.3f %12.4f %12.4f %12.4f" % ((n,) + tuple(r * 1e3 for r in res) + tuple(n / r / 1e6 for r in res)), flush=True)
