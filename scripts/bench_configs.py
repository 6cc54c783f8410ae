#!/usr/bin/env python3
"""BASELINE.json configs #3 and #4 on one B200 (device-resident inputs, CUDA events on the engine's stream):
  #3  2^k Groth16-shaped 4-pair pairing_batch checks: (a) flat (every G2 point walked on the fly),
      (b) three shared G2Precomp tables + one variable pair per check
  #4  2^20 G1 * Fr and 2^18 G2 * Fr
Prints one JSON line per measurement; parity of a sample against the oracle is asserted before timing."""
import argparse
import json
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import paritytech_bn_b200 as bn  # noqa: E402
from oracle import c_oracle as c  # noqa: E402
from tests import util  # noqa: E402

G1, G2, GT = 96, 192, 384
ap = argparse.ArgumentParser()
ap.add_argument("--checks-log2", type=int, default=16)
ap.add_argument("--g1-log2", type=int, default=20)
ap.add_argument("--g2-log2", type=int, default=18)
ap.add_argument("--reps", type=int, default=3)
args = ap.parse_args()

eng = bn.Engine(0)
dev = torch.device("cuda", 0)
stream = torch.cuda.ExternalStream(eng.stream, device=dev)


def timed(fn):
    best = None
    for _ in range(args.reps):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(stream); fn(); b.record(stream)
        torch.cuda.synchronize()
        t = a.elapsed_time(b)
        best = t if best is None else min(best, t)
    return best


def smul(which, ks):
    fr = util.fr_bytes(ks)
    if which == "g1":
        return eng.g1_mul(np.tile(util.G1_GEN, len(ks)), fr)
    return eng.g2_mul(np.tile(util.G2_GEN, len(ks)), fr)


# ---- config #3 ---------------------------------------------------------------------------------------------
n = 1 << args.checks_log2
w = util.groth16_checks(n, stream=91, scalar_mul=smul)
d_g1 = torch.from_numpy(w["g1"]).to(dev); d_g2 = torch.from_numpy(w["g2"]).to(dev)
d_out = torch.empty(n * GT, dtype=torch.uint8, device=dev); d_one = torch.empty(n, dtype=torch.uint8, device=dev)
eng.pairing_batch_dev(d_g1.data_ptr(), d_g2.data_ptr(), 4, n, d_out.data_ptr(), d_one.data_ptr()); eng.synchronize()
assert (d_one.cpu().numpy() == w["valid"]).all(), "accept/reject mismatch (flat)"
s = 64
exp = c.pairing_product_many(w["g1"][:s * 4 * G1], w["g2"][:s * 4 * G2], 4)[0]
assert d_out[:s * GT].cpu().numpy().tobytes() == exp.tobytes(), "Gt mismatch vs oracle (flat)"
t = timed(lambda: eng.pairing_batch_dev(d_g1.data_ptr(), d_g2.data_ptr(), 4, n, d_out.data_ptr(), d_one.data_ptr()))
print(json.dumps({"config": "#3 flat: 2^%d 4-pair pairing_batch checks" % args.checks_log2, "ms": t, "checks_per_s": n / (t * 1e-3),
                  "pairings_per_s": 4 * n / (t * 1e-3)}))
pc = eng.g2_precompute(w["vk_g2"])
g1s = w["g1"].reshape(n, 4, G1)[:, 1:4].reshape(-1).copy(); g1v = w["g1"].reshape(n, 4, G1)[:, 0].reshape(-1).copy()
g2v = w["g2"].reshape(n, 4, G2)[:, 0].reshape(-1).copy()
d_g1s = torch.from_numpy(g1s).to(dev); d_g1v = torch.from_numpy(g1v).to(dev); d_g2v = torch.from_numpy(g2v).to(dev)
d_out2 = torch.empty_like(d_out); d_one2 = torch.empty_like(d_one)
f = lambda: eng.pairing_batch_precomp_dev(pc, d_g1s.data_ptr(), d_g1v.data_ptr(), d_g2v.data_ptr(), 1, n, d_out2.data_ptr(), d_one2.data_ptr())  # noqa: E731
f(); eng.synchronize()
assert torch.equal(d_out, d_out2) and torch.equal(d_one, d_one2), "shared-table path differs from the flat path"
t = timed(f)
print(json.dumps({"config": "#3 shared G2Precomp (3 tables) + 1 variable pair: 2^%d checks" % args.checks_log2, "ms": t,
                  "checks_per_s": n / (t * 1e-3), "pairings_per_s": 4 * n / (t * 1e-3)}))
del d_g1, d_g2, d_out, d_one, d_out2, d_one2

# ---- config #4 ---------------------------------------------------------------------------------------------
for which, k, unit in (("g1", args.g1_log2, G1), ("g2", args.g2_log2, G2)):
    m = 1 << k
    base_k = util.scalars(1024, stream=95)
    base = smul(which, base_k).reshape(1024, unit)                        # Jacobian bases with z != 1
    bases = np.tile(base, (m // 1024, 1)).reshape(-1) if m >= 1024 else base[:m].reshape(-1)
    ks = util.scalars(min(m, 1 << 14), stream=96)
    ks[:4] = [0, 1, 2, util.R - 1]
    fr = np.tile(util.fr_bytes(ks).reshape(-1, 32), (max(1, m // len(ks)), 1)).reshape(-1)
    d_b = torch.from_numpy(bases).to(dev); d_k = torch.from_numpy(fr).to(dev); d_o = torch.empty(m * unit, dtype=torch.uint8, device=dev)
    fn = eng.g1_mul_dev if which == "g1" else eng.g2_mul_dev
    fn(d_b.data_ptr(), d_k.data_ptr(), d_o.data_ptr(), m); eng.synchronize()
    s = 256
    ref = (c.g1_mul if which == "g1" else c.g2_mul)(bases[:s * unit], fr[:s * 32])
    assert d_o[:s * unit].cpu().numpy().tobytes() == ref.tobytes(), "%s * Fr raw Jacobian mismatch vs oracle" % which
    t = timed(lambda: fn(d_b.data_ptr(), d_k.data_ptr(), d_o.data_ptr(), m))
    print(json.dumps({"config": "#4 2^%d %s * Fr" % (k, which.upper()), "ms": t, "muls_per_s": m / (t * 1e-3)}))
    del d_b, d_k, d_o
eng.close()
