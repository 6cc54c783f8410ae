"""Tiny driver for ncu: the quad-lane cooperative kernel at n = 1 (k_coop4_pairing) and Gt::pow at n = 1 (k_coop4_gt_pow)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402
import torch  # noqa: E402

import paritytech_bn_b200 as bn  # noqa: E402
from tests import util  # noqa: E402

eng = bn.Engine(0)
dev = torch.device("cuda", 0)
n = 8
g1 = eng.g1_mul(np.tile(util.G1_GEN, n), util.fr_images(n, stream=1))
g2 = eng.g2_mul(np.tile(util.G2_GEN, n), util.fr_images(n, stream=2))
d1, d2 = torch.from_numpy(g1).to(dev), torch.from_numpy(g2).to(dev)
out = torch.empty(n * 384, dtype=torch.uint8, device=dev)
torch.cuda.synchronize()
for m in (1, 1, 1):
    eng.pairing_dev(d1.data_ptr(), d2.data_ptr(), out.data_ptr(), m)
eng.synchronize()
e = out[:384].cpu().numpy()
eng.gt_pow(e, util.fr_images(1, stream=3))
print("done", eng.launch_count())
