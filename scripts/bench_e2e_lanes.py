"""End-to-end rate of bn254_pairing_many (pinned host buffers) against the number of pipeline lanes (BN254_PIPELINE = 0, 2, 3, 4)."""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402
import torch  # noqa: E402

import paritytech_bn_b200 as bn  # noqa: E402
from tests import util  # noqa: E402

k = int(sys.argv[1]) if len(sys.argv) > 1 else 16
n = 1 << k
base = 4096
e0 = bn.Engine(0)
p1 = e0.g1_mul(np.tile(util.G1_GEN, base), util.fr_images(base, stream=1)).reshape(base, 96)
p2 = e0.g2_mul(np.tile(util.G2_GEN, base), util.fr_images(base, stream=2)).reshape(base, 192)
idx = np.arange(n) % base
h1 = torch.from_numpy(p1[idx].reshape(-1).copy()).pin_memory()
h2 = torch.from_numpy(p2[(idx * 7 + 3) % base].reshape(-1).copy()).pin_memory()
ho = torch.empty(n * 384, dtype=torch.uint8, pin_memory=True)
ref = None
for lanes in (0, 2, 4, 6, 8, 4, 6, 8):
    e = bn.Engine(0, pipeline=lanes)
    e.pairing(h1.numpy(), h2.numpy(), out=ho.numpy())
    if ref is None:
        ref = ho.numpy().copy()
    assert ho.numpy().tobytes() == ref.tobytes()
    t0 = time.perf_counter()
    for _ in range(5):
        e.pairing(h1.numpy(), h2.numpy(), out=ho.numpy())
    dt = (time.perf_counter() - t0) / 5
    print("2^%d lanes %d: %.3f ms  %.3f M/s" % (k, lanes, dt * 1e3, n / dt / 1e6), flush=True)
    e.close()
