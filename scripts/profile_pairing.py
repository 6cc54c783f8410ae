"""Tiny driver for ncu: generate 2^k random pairs on the device, run the pairing path a few times."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import paritytech_bn_b200 as bn
from tests import util
k = int(sys.argv[1]) if len(sys.argv) > 1 else 16
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 2
variant = sys.argv[3] if len(sys.argv) > 3 else None
n = 1 << k
eng = bn.Engine(0, variant=variant)
dev = torch.device("cuda", 0)
fr1 = torch.from_numpy(util.fr_bytes(util.scalars(n, 1))).to(dev)
fr2 = torch.from_numpy(util.fr_bytes(util.scalars(n, 2))).to(dev)
gen1 = torch.from_numpy(np.tile(util.G1_GEN, n)).to(dev)
gen2 = torch.from_numpy(np.tile(util.G2_GEN, n)).to(dev)
g1 = torch.empty(n * 96, dtype=torch.uint8, device=dev)
g2 = torch.empty(n * 192, dtype=torch.uint8, device=dev)
out = torch.empty(n * 384, dtype=torch.uint8, device=dev)
torch.cuda.synchronize()
eng.g1_mul_dev(gen1.data_ptr(), fr1.data_ptr(), g1.data_ptr(), n)
eng.g2_mul_dev(gen2.data_ptr(), fr2.data_ptr(), g2.data_ptr(), n)
for _ in range(reps):
    eng.pairing_dev(g1.data_ptr(), g2.data_ptr(), out.data_ptr(), n)
eng.synchronize()
print("done", n, reps, eng.launch_count())
