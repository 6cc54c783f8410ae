"""Builds tests/abi/abi_smoke.c with gcc against include/bn254_b200.h and the shipped libbn254_b200.so: a compiled-language
consumer of the C ABI (what a Rust `bindgen`/`extern "C"` user meets), invoked from tests/test_gpu_parity.py."""
import os
import subprocess

_HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(_HERE))
EXE = os.path.join(_HERE, "_build", "abi_smoke")


def build():
    import paritytech_bn_b200 as bn
    lib = bn.library_path()
    src = os.path.join(_HERE, "abi_smoke.c")
    vec = os.path.join(_HERE, "golden_vectors.h")
    os.makedirs(os.path.dirname(EXE), exist_ok=True)
    deps = [src, vec, os.path.join(ROOT, "include", "bn254_b200.h"), lib]
    if not os.path.exists(EXE) or os.path.getmtime(EXE) < max(os.path.getmtime(d) for d in deps):
        tmp = "%s.tmp.%d" % (EXE, os.getpid())
        subprocess.check_call(["/usr/bin/gcc", "-std=c11", "-O1", "-Wall", "-Wextra", "-Werror", "-I", os.path.join(ROOT, "include"),
                               "-o", tmp, src, "-L", os.path.dirname(lib), "-lbn254_b200", "-Wl,-rpath," + os.path.dirname(lib)])
        os.replace(tmp, EXE)
    return EXE
