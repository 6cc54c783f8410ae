#!/usr/bin/env python3
"""torchrun payload (one rank per GPU, NCCL): ONE pairing_batch product sharded over the ranks with the 384-byte
partial gather (paritytech-bn_b200/distributed.py), compared on every rank with the single-GPU pairing_batch of the
whole set and with the known answer (the scalars are chosen so that the product is one, then perturbed)."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import paritytech_bn_b200 as bn  # noqa: E402
from oracle import bn254_ref as o  # noqa: E402
from tests import util  # noqa: E402

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
eng = bn.Engine(local)
n = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
ks1, ks2 = util.scalars(n, stream=101), util.scalars(n, stream=102)
tot = sum(x * y for x, y in zip(ks1[:-1], ks2[:-1])) % o.R_ORDER
ks1[-1] = (-tot * pow(ks2[-1], -1, o.R_ORDER)) % o.R_ORDER          # sum a_i b_i = 0 mod r: the product is one
g1 = eng.g1_mul(np.tile(util.G1_GEN, n), util.fr_bytes(ks1))
g2 = eng.g2_mul(np.tile(util.G2_GEN, n), util.fr_bytes(ks2))
lo, hi = bn.shard_range(n, rank, world)
out, is_one = bn.pairing_batch_distributed(eng, g1[lo * 96:hi * 96], g2[lo * 192:hi * 192], device=torch.device("cuda", local))
assert is_one and out.tobytes() == util.GT_ONE.tobytes(), "sharded product is not one"
g1[:96] = util.G1_GEN                                                # perturb: no longer one
out, is_one = bn.pairing_batch_distributed(eng, g1[lo * 96:hi * 96], g2[lo * 192:hi * 192], device=torch.device("cuda", local))
ref, ref_one = eng.pairing_batch(g1, g2)
assert not is_one and out.tobytes() == ref.tobytes(), "sharded product differs from the single-GPU pairing_batch"
# independent pairings: every rank computes its own index range, rank 0 checks the concatenation
mine = eng.pairing(g1[lo * 96:hi * 96], g2[lo * 192:hi * 192])
if rank == 0:
    whole = eng.pairing(g1, g2)
    assert whole[lo * 384:hi * 384].tobytes() == mine.tobytes()
dist.barrier()
print("rank %d/%d ok: sharded product == single-GPU product, n = %d" % (rank, world, n), flush=True)
dist.destroy_process_group()
eng.close()
