"""Loader for the TEST-ONLY host build of the device headers (see hostsim.cpp)."""
import ctypes
import glob
import os
import subprocess

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "_build", "libhostsim.so")
_CSRC = os.path.join(_HERE, "..", "..", "paritytech-bn_b200", "csrc")


def build():
    srcs = [os.path.join(_HERE, "hostsim.cpp")] + glob.glob(os.path.join(_CSRC, "*.cuh"))
    if not os.path.exists(_SO) or os.path.getmtime(_SO) < max(os.path.getmtime(s) for s in srcs):
        os.makedirs(os.path.dirname(_SO), exist_ok=True)
        subprocess.check_call(["/usr/bin/g++", "-O2", "-shared", "-fPIC", "-Wno-unknown-pragmas", "-o", _SO,
                               os.path.join(_HERE, "hostsim.cpp")])
    return _SO


_lib = None


def lib():
    global _lib
    if _lib is None:
        _lib = ctypes.CDLL(build())
    return _lib
