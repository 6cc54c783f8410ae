"""Tiny driver for ncu: 2^18 G1 and 2^16 G2 scalar multiplications, reference-sequence kernel and window kernel, twice each."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402
import torch  # noqa: E402

import paritytech_bn_b200 as bn  # noqa: E402
from tests import util  # noqa: E402

eng = bn.Engine(0)
dev = torch.device("cuda", 0)
for which, k, unit in (("g1", 18, 96), ("g2", 16, 192)):
    n = 1 << k
    gen = torch.from_numpy(np.tile(util.G1_GEN if which == "g1" else util.G2_GEN, n)).to(dev)
    fr0 = torch.from_numpy(util.fr_images(n, stream=1)).to(dev)
    fr = torch.from_numpy(util.fr_images(n, stream=2)).to(dev)
    base = torch.empty(n * unit, dtype=torch.uint8, device=dev)
    out = torch.empty(n * unit, dtype=torch.uint8, device=dev)
    torch.cuda.synchronize()
    mul, muln = (eng.g1_mul_dev, eng.g1_mul_normalized_dev) if which == "g1" else (eng.g2_mul_dev, eng.g2_mul_normalized_dev)
    mul(gen.data_ptr(), fr0.data_ptr(), base.data_ptr(), n)          # launch 1 (g1) / 5 (g2): bases
    mul(base.data_ptr(), fr.data_ptr(), out.data_ptr(), n)
    muln(base.data_ptr(), fr.data_ptr(), out.data_ptr(), n)
    muln(base.data_ptr(), fr.data_ptr(), out.data_ptr(), n)
    eng.synchronize()
print("done", eng.launch_count())
