"""Per-kernel milliseconds of the 2^16 independent-pairing step for several builds of the library (A/B of kernel experiments):
python scripts/ab_variants.py "" _hoist _pf ...   (variant suffixes of paritytech-bn_b200/lib/libbn254_b200<suffix>.so)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402
import torch  # noqa: E402

import paritytech_bn_b200 as bn  # noqa: E402
from tests import util  # noqa: E402

variants = sys.argv[1:] or [""]
n, base = 1 << 16, 4096
dev = torch.device("cuda", 0)
e0 = bn.Engine(0)
p1 = e0.g1_mul(np.tile(util.G1_GEN, base), util.fr_images(base, stream=1)).reshape(base, 96)
p2 = e0.g2_mul(np.tile(util.G2_GEN, base), util.fr_images(base, stream=2)).reshape(base, 192)
idx = np.arange(n) % base
d1 = torch.from_numpy(p1[idx].reshape(-1).copy()).to(dev)
d2 = torch.from_numpy(p2[(idx * 7 + 3) % base].reshape(-1).copy()).to(dev)
ref = None
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
for rnd in range(2):
    for v in variants:
        e = bn.Engine(0, variant=v)
        do = torch.empty(n * 384, dtype=torch.uint8, device=dev)
        e.pairing_timed_dev(d1.data_ptr(), d2.data_ptr(), do.data_ptr(), n)
        got = do.cpu().numpy()
        if ref is None:
            ref = got
        assert got.tobytes() == ref.tobytes(), v
        rows = []
        for _ in range(7):
            flush.fill_(1)
            torch.cuda.synchronize()
            rows.append(e.pairing_timed_dev(d1.data_ptr(), d2.data_ptr(), do.data_ptr(), n))
        med = np.median(np.array(rows), axis=0)
        print("%-8s affine %.3f lines %.3f miller %.3f easy %.3f hard %.3f  total %.3f ms" % ((v or "base",) + tuple(med) + (med.sum(),)), flush=True)
        e.close()
