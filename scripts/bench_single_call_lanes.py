"""bn254_pairing_many_sharded from ONE process over all visible GPUs (2^17 pairs per GPU, pinned host buffers) against the lane
count of the host-buffer pipeline (option pipeline = 0, 2, 4)."""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402
import torch  # noqa: E402

import paritytech_bn_b200 as bn  # noqa: E402
from tests import util  # noqa: E402

world = torch.cuda.device_count()
n = world << 17
base = 1 << 14
e0 = bn.Engine(0)
p1 = e0.g1_mul(np.tile(util.G1_GEN, base), util.fr_images(base, stream=501)).reshape(base, 96)
p2 = e0.g2_mul(np.tile(util.G2_GEN, base), util.fr_images(base, stream=502)).reshape(base, 192)
e0.close()
idx = (np.arange(n, dtype=np.int64) * 2654435761 % base)
h1 = torch.from_numpy(p1[idx].reshape(-1)).pin_memory()
h2 = torch.from_numpy(p2[(idx * 7 + 3) % base].reshape(-1)).pin_memory()
ho = torch.empty(n * 384, dtype=torch.uint8, pin_memory=True)
ref = None
for lanes in (2, 4, 0, 2, 4):
    engines = [bn.Engine(d, pipeline=lanes) for d in range(world)]
    bn.pairing_sharded(engines, h1.numpy(), h2.numpy(), out=ho.numpy())
    if ref is None:
        ref = ho.numpy()[:4096 * 384].copy()
    assert ho.numpy()[:4096 * 384].tobytes() == ref.tobytes()
    ts = []
    for _ in range(5):
        t0 = time.perf_counter()
        bn.pairing_sharded(engines, h1.numpy(), h2.numpy(), out=ho.numpy())
        ts.append(time.perf_counter() - t0)
    print("devices %d lanes %d: median %.2f ms  min %.2f ms  %.2f M/s" % (world, lanes, 1e3 * sorted(ts)[2], 1e3 * min(ts), n / sorted(ts)[2] / 1e6), flush=True)
    for e in engines:
        e.close()
