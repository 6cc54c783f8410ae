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
                 "bn254_pairing_product_precomp_many", "bn254_pairing_product_sharded", "bn254_pairing_many_sharded",
                 "bn254_ctx_create_ex", "bn254_ctx_options_default", "bn254_g1_mul_normalized_many", "bn254_g2_mul_normalized_many",
                 "bn254_miller_product_dev"):
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


def test_header_compiles_as_c_and_the_c_consumer_links():
    """include/bn254_b200.h is plain C11 (gcc -Wall -Wextra -Werror) and tests/abi/abi_smoke.c links against the shipped
    library through it alone (the program itself runs in the GPU suite)."""
    from tests import abi
    exe = abi.build()
    assert os.path.exists(exe)


def test_options_struct_mirror_matches_the_header():
    """The ctypes mirror of bn254_ctx_options has the header's field list in the header's order."""
    import paritytech_bn_b200 as bn
    src = open(os.path.join(ROOT, "include", "bn254_b200.h")).read()
    body = src[src.index("typedef struct bn254_ctx_options {"):src.index("} bn254_ctx_options;")]
    fields = re.findall(r"^\s*(?:uint32_t|int32_t)\s+([a-z_0-9]+);", body, flags=re.M)
    assert fields == [f for f, _ in bn.CtxOptions._fields_]
    o = bn.CtxOptions(fexp_smem=0)
    assert o.struct_size == 4 * len(fields) and o.fexp_smem == 0 and o.pipeline == -1


def test_limb_arrays_are_reinterpreted_not_value_cast():
    """a uint64/uint32 limb array (the natural U256 layout) must reach the library as its bytes."""
    import numpy as np
    from paritytech_bn_b200.api import _arr
    limbs = np.array([0x0123456789abcdef, 2, 3, 4], dtype=np.uint64)
    b, n = _arr(limbs, 32)
    assert n == 1 and b.tobytes() == limbs.tobytes()
    b32, n32 = _arr(limbs.view(np.uint32), 32)
    assert n32 == 1 and b32.tobytes() == limbs.tobytes()
    with pytest.raises(TypeError):
        _arr(np.zeros(4, dtype=np.float64), 32)
    with pytest.raises(ValueError):
        _arr(np.zeros(5, dtype=np.uint64), 32)
