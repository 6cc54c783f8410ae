"""Tiny driver for ncu: 2^16 Groth16-shaped 4-pair checks (BASELINE configs[2]), flat (pairing_batch: all four line tables built)
and with the three shared G2Precomp tables (pairing_batch_precomp), twice each; results checked against the construction."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402

import paritytech_bn_b200 as bn  # noqa: E402
from tests import util  # noqa: E402

eng = bn.Engine(0)
n = 1 << 16


def smul(which, ks):
    fr = util.fr_bytes(ks)
    if which == "g1":
        return eng.g1_mul(np.tile(util.G1_GEN, len(ks)), fr)
    return eng.g2_mul(np.tile(util.G2_GEN, len(ks)), fr)


w = util.groth16_checks(n, stream=7, scalar_mul=smul)
pc = eng.g2_precompute(w["vk_g2"])
for _ in range(2):
    out, is_one = eng.pairing_batch(w["g1"], w["g2"], pairs_per_check=4)
    assert (is_one == w["valid"]).all()
    out2, is_one2 = eng.pairing_batch_precomp(pc, w["g1_shared"], w["g1_var"], w["g2_var"], 1)
    assert (is_one2 == w["valid"]).all() and out2.tobytes() == out.tobytes()
pc.close()
print("done", eng.launch_count())
