"""CPU-side checks of the drop-in boundary: the C-ABI library loads and exports every symbol that
include/bn254_b200.h declares, and refuses to run without a GPU (no CPU fallback)."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    src = open(os.path.join(ROOT, "include", "bn254_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(bn254_[a-z0-9_]+)\s*\(", src)))


def test_header_declares_the_surface():
    names = _declared()
    for must in ("bn254_pairing_many", "bn254_pairing_product_many", "bn254_miller_loop_batch", "bn254_final_exp_many",
                 "bn254_gt_eq_many", "bn254_g1_mul_many", "bn254_g2_mul_many", "bn254_g2_precompute",
                 "bn254_pairing_product_precomp_many", "bn254_pairing_product_sharded"):
        assert must in names


@pytest.mark.parametrize("variant", [""])
def test_library_exports_every_declared_symbol(variant):
    import paritytech_bn_b200 as bn
    if not os.path.exists(bn.library_path(variant)):
        import __graft_entry__
        __graft_entry__.build()
    lib = ctypes.CDLL(bn.library_path(variant))
    missing = [n for n in _declared() if not hasattr(lib, n)]
    assert not missing, missing


def test_no_cpu_fallback():
    import torch

    import paritytech_bn_b200 as bn
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(bn.BnError):
        bn.Engine(0)
