#!/usr/bin/env python3
"""SURVEY 8f N2/N3 throughput on one B200, through the C ABI with HOST buffers (copies inside the timed region,
wall clock around the call): AffineG::new validation and from_compressed decoding, with the C oracle (OpenMP, all
host cores) timed on a bounded sample beside each.  One JSON line per measurement; parity of the sample against
the oracle is asserted before timing."""
import argparse
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import paritytech_bn_b200 as bn  # noqa: E402
from oracle import bn254_ref as o  # noqa: E402
from oracle import c_oracle as c  # noqa: E402
from tests import util  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--g2-log2", type=int, default=16)
ap.add_argument("--g1-log2", type=int, default=20)
ap.add_argument("--reps", type=int, default=3)
ap.add_argument("--cpu-sample", type=int, default=2048)
ap.add_argument("--tag", default=os.environ.get("BN254_SUBGROUP_REF", "0") == "1" and "reference ladder" or "psi test")
args = ap.parse_args()
eng = bn.Engine(0)


def best(fn):
    t = None
    for _ in range(args.reps):
        t0 = time.perf_counter(); fn(); dt = time.perf_counter() - t0
        t = dt if t is None else min(t, dt)
    return t


def points(n1, n2):
    k1, k2 = util.scalars(n1, stream=61), util.scalars(n2, stream=62)
    g1 = eng.g1_normalize(eng.g1_mul(np.tile(util.G1_GEN, n1), util.fr_bytes(k1)))
    g2 = eng.g2_normalize(eng.g2_mul(np.tile(util.G2_GEN, n2), util.fr_bytes(k2)))
    return g1, g2


def compress(g1, g2):
    """Vectorised host-side compression of normalised Montgomery images (big-int work only for the tags)."""
    e1, e2 = [], []
    for row in g1.reshape(-1, 96):
        x, y, _ = o.g1_from_bytes(row.tobytes())
        e1.append(util.compress_g1(x, y))
    for row in g2.reshape(-1, 192):
        x, y, _ = o.g2_from_bytes(row.tobytes())
        e2.append(util.compress_g2(x, y))
    return np.frombuffer(b"".join(e1), dtype=np.uint8).copy(), np.frombuffer(b"".join(e2), dtype=np.uint8).copy()


def line(metric, n, t_gpu, cpu_n, t_cpu):
    print(json.dumps({"metric": metric, "value": n / t_gpu, "unit": "points/s", "n": n, "ms": t_gpu * 1e3,
                      "g2_order_check": args.tag, "timing": "wall clock around the C-ABI call, host buffers",
                      "cpu_baseline": {"value": cpu_n / t_cpu, "unit": "points/s", "cores": c.num_procs(), "kind": "port",
                                       "sample": "%d points" % cpu_n}}), flush=True)


n1, n2 = 1 << args.g1_log2, 1 << args.g2_log2
g1, g2 = points(n1, n2)
a1, a2 = util.affine_of(g1, 96), util.affine_of(g2, 192)
a2.reshape(-1, 128)[::5, 3] ^= 2            # every fifth point off the curve
s = args.cpu_sample

# ---- N2: AffineG::new ----
st = eng.g2_affine_check(a2)
t0 = time.perf_counter(); ref = c.g2_affine_check(a2[:s * 128]); t_cpu = time.perf_counter() - t0
assert st[:s].tobytes() == ref.tobytes() and set(st) == {0, 1}
line("g2_affine_new_per_s", n2, best(lambda: eng.g2_affine_check(a2)), s, t_cpu)
st = eng.g1_affine_check(a1)
t0 = time.perf_counter(); ref = c.g1_affine_check(a1); t_cpu = time.perf_counter() - t0
assert st.tobytes() == ref.tobytes() and not st.any()
line("g1_affine_new_per_s", n1, best(lambda: eng.g1_affine_check(a1)), n1, t_cpu)

# ---- N3: from_compressed ----
m1, m2 = min(n1, 1 << 16), min(n2, 1 << 14)  # host-side compression is pure Python: bound it, then tile
e1, e2 = compress(g1[:m1 * 96], g2[:m2 * 192])
e1, e2 = np.tile(e1, n1 // m1), np.tile(e2, n2 // m2)
out, st = eng.g2_from_compressed(e2)
t0 = time.perf_counter(); ro, so = c.g2_from_compressed(e2[:s * 65]); t_cpu = time.perf_counter() - t0
assert not st.any() and out[:s * 192].tobytes() == ro.tobytes() and out[:m2 * 192].tobytes() == g2[:m2 * 192].tobytes()
line("g2_from_compressed_per_s", n2, best(lambda: eng.g2_from_compressed(e2)), s, t_cpu)
out, st = eng.g1_from_compressed(e1)
sc = min(n1, 1 << 16)
t0 = time.perf_counter(); ro, so = c.g1_from_compressed(e1[:sc * 33]); t_cpu = time.perf_counter() - t0
assert not st.any() and out[:sc * 96].tobytes() == ro.tobytes() and out[:m1 * 96].tobytes() == g1[:m1 * 96].tobytes()
line("g1_from_compressed_per_s", n1, best(lambda: eng.g1_from_compressed(e1)), sc, t_cpu)
