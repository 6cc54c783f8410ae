#!/usr/bin/env python3
"""Every host-buffer C-ABI entry point once, at a ragged size (not a multiple of the 64-thread block), each result
compared with the oracle.  Written as the payload for compute-sanitizer; that tool is closed on this GPU pool, so
the script runs plain (tests/test_gpu_wire.py::test_every_entry_point_at_a_ragged_size)."""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import paritytech_bn_b200 as bn  # noqa: E402
from oracle import c_oracle as c  # noqa: E402
from tests import util  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 70
eng = bn.Engine(0)
ks = util.scalars(n, stream=81)
g1, g2 = util.points_cpu(ks, ks[::-1])
g1[:96] = util.G1_INF                                         # one pair with a point at infinity
fr = util.fr_bytes(util.scalars(n, stream=82))

assert eng.pairing(g1, g2).tobytes() == c.pairing_many(g1, g2).tobytes()
out, one = eng.pairing_batch(g1[: (n // 3) * 3 * 96], g2[: (n // 3) * 3 * 192], 3)
exp, eone = c.pairing_product_many(g1[: (n // 3) * 3 * 96], g2[: (n // 3) * 3 * 192], 3)
assert out.tobytes() == exp.tobytes()
out, one = eng.pairing_batch(g1[: (n // 9) * 9 * 96], g2[: (n // 9) * 9 * 192], 9)   # > 8 pairs: independent loops + group product
exp, eone = c.pairing_product_many(g1[: (n // 9) * 9 * 96], g2[: (n // 9) * 9 * 192], 9)
assert out.tobytes() == exp.tobytes()
m, ok = eng.miller_loop_many(g2, g1)
mo, oko = c.miller_loop_many(g2, g1)
assert m.tobytes() == mo.tobytes() and ok.tobytes() == oko.tobytes()
fe, some = eng.final_exponentiation(m)
assert fe.tobytes() == c.final_exp(m)[0].tobytes()
assert eng.miller_product(g1, g2).tobytes() == c.miller_loop_batch_skip(g2, g1).tobytes()
assert eng.gt_mul(fe, fe[::1]).tobytes() == c.fq12_mul(fe, fe).tobytes()
assert eng.gt_pow(fe, fr).tobytes() == c.gt_pow(fe, fr).tobytes()
assert eng.gt_inverse(fe)[0].tobytes() == c.fq12_inverse(fe)[0].tobytes()
assert eng.gt_eq(fe, fe).all()
p1, p2 = eng.g1_mul(g1, fr), eng.g2_mul(g2, fr)
assert p1.tobytes() == c.g1_mul(g1, fr).tobytes() and p2.tobytes() == c.g2_mul(g2, fr).tobytes()
n1, n2 = eng.g1_normalize(p1), eng.g2_normalize(p2)
assert n1.tobytes() == c.g1_normalize(p1).tobytes() and n2.tobytes() == c.g2_normalize(p2).tobytes()
a1, a2 = util.affine_of(n1[96:], 96), util.affine_of(n2, 192)
a2.reshape(-1, 128)[::4, 9] ^= 1
assert eng.g1_affine_check(a1).tobytes() == c.g1_affine_check(a1).tobytes()
assert eng.g2_affine_check(a2).tobytes() == c.g2_affine_check(a2).tobytes()
from tests.test_wire_oracle import wire_cases_g1, wire_cases_g2  # noqa: E402
for cases, fn, ref in ((wire_cases_g1(), eng.g1_from_compressed, c.g1_from_compressed), (wire_cases_g2(), eng.g2_from_compressed, c.g2_from_compressed)):
    enc = b"".join(e for e, _ in cases)
    o1, s1 = fn(enc)
    o2, s2 = ref(enc)
    assert o1.tobytes() == o2.tobytes() and s1.tobytes() == s2.tobytes()
# shared G2Precomp tables: 3 shared pairs + 1 variable pair per check
w = util.groth16_checks(n, stream=83)
out, one = eng.pairing_batch(w["g1"], w["g2"], 4)
assert (one == w["valid"]).all()
pre = eng.g2_precompute(w["vk_g2"])
out2, one2 = eng.pairing_batch_precomp(pre, w["g1_shared"], w["g1_var"], w["g2_var"], 1)
assert out2.tobytes() == out.tobytes() and (one2 == one).all()
assert pre.export(0).tobytes() == c.g2_precompute(w["vk_g2"][:192]).tobytes()
pre.close()
eng.close()
print("all_entry_points: every entry point ran, results match the oracle, n =", n)
