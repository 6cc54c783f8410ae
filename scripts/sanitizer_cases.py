"""Small invocations of every shared-memory kernel for compute-sanitizer (memcheck / racecheck / synccheck):
smoke()'s calls plus the n = 449 and n = 1600 cases of test_slot_file_and_local_memory_paths_agree, a product check on
512-thread blocks and the window scalar multiplication.  Results are compared with the oracle, so a race that corrupts
data fails here even if the tool misses it.  usage: compute-sanitizer --tool racecheck python scripts/sanitizer_cases.py"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402

import paritytech_bn_b200 as bn  # noqa: E402
from oracle import c_oracle as c  # noqa: E402
from tests import util  # noqa: E402

sizes = [int(x) for x in sys.argv[1:]] or [8, 449, 1600]
eng = bn.Engine(0, coop_max_n=0)          # the slot-file kernels (one thread per pairing) at small sizes
coop = bn.Engine(0, coop4_max_n=0)        # one warp per pairing (k_coop_pairing)
quad = bn.Engine(0)                       # default: batches of at most 592 run four warps per pairing (k_coop4_pairing), larger ones one warp
for n in sizes:
    g1, g2 = util.points_cpu(util.scalars(n, stream=71), util.scalars(n, stream=72))
    g1 = g1.reshape(n, 96).copy()
    g1[min(3, n - 1)] = util.G1_INF
    g1 = g1.reshape(-1)
    m = min(n, 64)
    exp = c.pairing_many(g1[:m * 96], g2[:m * 192]).tobytes()
    for name, e in (("slot-file", eng), ("warp-cooperative", coop), ("quad-lane / default", quad)):
        got = e.pairing(g1, g2)
        assert got[:m * 384].tobytes() == exp, (name, n)
        print("pairing n=%d %s ok" % (n, name), flush=True)
    if n <= 8:
        ks = util.fr_bytes(util.scalars(n, stream=76))
        want = c.gt_pow(got, ks).tobytes()
        for name, e in (("warp-cooperative", coop), ("quad-lane", quad)):
            assert e.gt_pow(got, ks).tobytes() == want, (name, n)
            ml, _ = e.miller_loop_many(g2, g1)
            fe, _ = e.final_exponentiation(ml)
            assert fe.tobytes() == got.tobytes(), (name, n)
        print("gt_pow / miller / final exp n=%d on both cooperative executors ok" % n, flush=True)
coop.close()
quad.close()
w = util.groth16_checks(70)
out, is_one = eng.pairing_batch(w["g1"], w["g2"], pairs_per_check=4)
rout, ris = c.pairing_product_many(w["g1"], w["g2"], 4)
assert out.tobytes() == rout.tobytes() and (is_one == ris).all()
pc = eng.g2_precompute(w["vk_g2"])
out2, _ = eng.pairing_batch_precomp(pc, w["g1_shared"], w["g1_var"], w["g2_var"], 1)
assert out2.tobytes() == rout.tobytes()
pc.close()
print("product checks ok", flush=True)
e2 = bn.Engine(0, big_blocks=2)
n = 600
g1, g2 = util.points_cpu(util.scalars(n, stream=73), util.scalars(n, stream=74))
assert e2.pairing(g1, g2)[:32 * 384].tobytes() == c.pairing_many(g1[:32 * 96], g2[:32 * 192]).tobytes()
o3, _ = e2.pairing_batch(g1, g2, pairs_per_check=1)
assert o3[:32 * 384].tobytes() == c.pairing_many(g1[:32 * 96], g2[:32 * 192]).tobytes()
e2.close()
print("512-thread blocks ok", flush=True)
fr = util.fr_bytes(util.scalars(100, stream=75))
assert eng.g1_mul_normalized(g1[:100 * 96], fr).tobytes() == c.g1_normalize(c.g1_mul(g1[:100 * 96], fr)).tobytes()
assert eng.g2_mul_normalized(g2[:100 * 192], fr).tobytes() == c.g2_normalize(c.g2_mul(g2[:100 * 192], fr)).tobytes()
e3 = bn.Engine(0)
outs, one = bn.pairing_product_sharded([eng, e3], g1[:40 * 96], g2[:40 * 192])
assert outs.tobytes() == c.pairing_batch(g1[:40 * 96], g2[:40 * 192]).tobytes()
got = bn.pairing_sharded([eng, e3], g1[:70 * 96], g2[:70 * 192])
assert got.tobytes() == c.pairing_many(g1[:70 * 96], g2[:70 * 192]).tobytes()
e3.close()
print("scalar mul / sharded ok", flush=True)
eng.close()
print("sanitizer cases ok")
