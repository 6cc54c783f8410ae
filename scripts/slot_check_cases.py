"""Stand-in for compute-sanitizer (closed on this pool): runs the small cases of every shared-memory kernel on the
-DBN_SLOT_CHECK build (libbn254_b200_chk.so), whose slot accessors verify that each access stays inside the memory its issuer
owns (bn_slots.cuh), compares every result with the oracle, and reads the violation counter: it must be 0 after the cases and
exactly 1 after the deliberate out-of-region store of the negative control.
usage: python scripts/slot_check_cases.py   (build first: make -C paritytech-bn_b200/csrc SUFFIX=_chk EXTRA=-DBN_SLOT_CHECK)"""
import ctypes
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
os.environ["BN254_B200_VARIANT"] = "_chk"
import paritytech_bn_b200 as bn  # noqa: E402


def counter(eng):
    out = (ctypes.c_ulonglong * 4)()
    eng._check(eng.lib.bn254_debug_slot_violations(eng.ctx, out))
    return [int(x) for x in out]


eng = bn.Engine(0)
print("library:", bn.library_path(), flush=True)
eng._check(eng.lib.bn254_debug_slot_reset(eng.ctx))
assert counter(eng)[0] == 0
# the cases: pairings on the slot-file kernels, the warp-cooperative and the quad-lane kernels at n = 8, 449, 1600 (+ Gt::pow at 8); product checks flat and
# with shared tables; 512-thread blocks on 6 slots; window scalar multiplication; sharded product / pairing over two contexts
# (run in THIS process: the counter is a device global of the library image loaded here)
import runpy  # noqa: E402
sys.argv = ["sanitizer_cases.py", "8", "449", "1600"]
runpy.run_path(os.path.join(ROOT, "scripts", "sanitizer_cases.py"), run_name="__main__")
c = counter(eng)
print("violations after the cases:", c, flush=True)
assert c[0] == 0, "shared-memory access outside the issuer's region: %s" % c
eng._check(eng.lib.bn254_debug_slot_poke(eng.ctx))
c = counter(eng)
print("violations after the negative control (thread 5 stores into thread 6's slots):", c, flush=True)
assert c[0] == 1 and (c[1] >> 32) == 1 and (c[3] & 0xffffffff) == 5
print("slot check ok: 0 violations in the cases, the negative control is caught")
