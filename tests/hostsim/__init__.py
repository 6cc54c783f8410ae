"""Loader for the TEST-ONLY host build of the device headers (see hostsim.cpp)."""
import ctypes
import glob
import os
import subprocess

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "_build", "libhostsim.so")
_CSRC = os.path.join(_HERE, "..", "..", "paritytech-bn_b200", "csrc")


def build():
    """Lock + atomic rename: safe when several test processes start on a fresh checkout."""
    import fcntl
    srcs = [os.path.join(_HERE, "hostsim.cpp")] + glob.glob(os.path.join(_CSRC, "*.cuh"))
    os.makedirs(os.path.dirname(_SO), exist_ok=True)
    with open(os.path.join(os.path.dirname(_SO), ".lock"), "w") as lk:
        fcntl.flock(lk, fcntl.LOCK_EX)
        if not os.path.exists(_SO) or os.path.getmtime(_SO) < max(os.path.getmtime(s) for s in srcs):
            tmp = "%s.tmp.%d" % (_SO, os.getpid())
            subprocess.check_call(["/usr/bin/g++", "-O2", "-shared", "-fPIC", "-Wno-unknown-pragmas", "-o", tmp,
                                   os.path.join(_HERE, "hostsim.cpp")])
            os.replace(tmp, _SO)
    return _SO


_lib = None


def lib():
    global _lib
    if _lib is None:
        _lib = ctypes.CDLL(build())
    return _lib
