"""ctypes loader for the C oracle (oracle/bn254_oracle.c).  TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import this.
All buffers are numpy uint8 arrays holding the crate's repr(C) Montgomery images.
"""
import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "_build", "libbn254_oracle.so")
# bench.py's CPU legs set BN254_ORACLE_NATIVE=1: the library is then compiled ON the machine that times it with -march=native
#  instead of the portable x86-64-v3 build that travels with the snapshot.
_SO_NATIVE = os.path.join(_HERE, "_build", "libbn254_oracle_native.so")
NATIVE_CFLAGS = "-O3 -march=native -fopenmp -fPIC -Wall -Wextra -Wno-unused-function"

FQ, FR, FQ2, FQ6, FQ12 = 32, 32, 64, 192, 384
G1, G2, ELL = 96, 192, 192
G2PRECOMP = 128 + 87 * 192


def build(force=False, native=False):
    """Compile the oracle if missing or stale.  Serialised with a lock file and installed by an atomic rename, so
    concurrent processes (the two ranks of the gloo test on a fresh checkout) never load a half-written library.
    native=True: a second library built with -march=native on this very machine (always rebuilt when the host differs:
    the marker file records the CPU model it was built on)."""
    import fcntl
    src = os.path.join(_HERE, "bn254_oracle.c")
    so = _SO_NATIVE if native else _SO
    os.makedirs(os.path.dirname(so), exist_ok=True)
    with open(os.path.join(os.path.dirname(so), ".lock"), "w") as lk:
        fcntl.flock(lk, fcntl.LOCK_EX)
        stale = force or not os.path.exists(so) or os.path.getmtime(so) < os.path.getmtime(src)
        marker, cpu = so + ".cpu", ""
        if native:
            try:
                cpu = [ln for ln in open("/proc/cpuinfo") if ln.startswith(("model name", "flags"))][:2]
                cpu = "".join(cpu)
            except OSError:
                pass
            stale = stale or not os.path.exists(marker) or open(marker).read() != cpu
        if stale:
            tmp = "%s.tmp.%d" % (so, os.getpid())
            cmd = ["make", "-s", "-B", "-C", _HERE, "OUT=" + tmp]
            if native:
                cmd.append("CFLAGS=" + NATIVE_CFLAGS)
            subprocess.check_call(cmd)
            os.replace(tmp, so)
            if native:
                with open(marker, "w") as f:
                    f.write(cpu)
    return so


_lib = None


def lib():
    global _lib
    if _lib is None:
        so = None
        if os.environ.get("BN254_ORACLE_NATIVE") == "1":
            try:
                so = build(native=True)
            except Exception:
                so = None          # no compiler on this machine: the portable build below
        if so is None:
            so = build()
        _lib = ctypes.CDLL(so)
        _lib.build_flavour = "native" if so == _SO_NATIVE else "x86-64-v3"
        _lib.orc_init()
        _lib.orc_g2precomp_size.restype = ctypes.c_size_t
        assert _lib.orc_g2precomp_size() == G2PRECOMP
    return _lib


def _p(a):
    return a.ctypes.data_as(ctypes.c_void_p)


def _in(a, unit):
    a = np.ascontiguousarray(np.frombuffer(a, dtype=np.uint8) if isinstance(a, (bytes, bytearray)) else a, dtype=np.uint8).reshape(-1)
    assert a.size % unit == 0, (a.size, unit)
    return a, a.size // unit


def num_procs():
    return lib().orc_num_procs()


def _binary(name, unit):
    def f(a, b):
        a, n = _in(a, unit)
        b, nb = _in(b, unit)
        assert n == nb
        out = np.empty(n * unit, dtype=np.uint8)
        getattr(lib(), name)(_p(a), _p(b), _p(out), ctypes.c_size_t(n))
        return out
    return f


def _unary(name, unit):
    def f(a):
        a, n = _in(a, unit)
        out = np.empty(n * unit, dtype=np.uint8)
        getattr(lib(), name)(_p(a), _p(out), ctypes.c_size_t(n))
        return out
    return f


fq_mul = _binary("orc_fq_mul_many", FQ)
fq_add = _binary("orc_fq_add_many", FQ)
fq_sub = _binary("orc_fq_sub_many", FQ)
fq2_mul = _binary("orc_fq2_mul_many", FQ2)
fq2_sqr = _unary("orc_fq2_sqr_many", FQ2)
fq6_mul = _binary("orc_fq6_mul_many", FQ6)
fq12_mul = _binary("orc_fq12_mul_many", FQ12)
fq12_add = _binary("orc_fq12_add_many", FQ12)
fq12_sub = _binary("orc_fq12_sub_many", FQ12)
fq12_sqr = _unary("orc_fq12_sqr_many", FQ12)
fq12_neg = _unary("orc_fq12_neg_many", FQ12)
fq12_cyclotomic_squared = _unary("orc_fq12_cyclotomic_squared_many", FQ12)
fq12_exp_by_neg_z = _unary("orc_fq12_exp_by_neg_z_many", FQ12)
g1_add = _binary("orc_g1_add_many", G1)
g2_add = _binary("orc_g2_add_many", G2)
g1_double = _unary("orc_g1_double_many", G1)
g2_double = _unary("orc_g2_double_many", G2)
g1_normalize = _unary("orc_g1_normalize_many", G1)
g2_normalize = _unary("orc_g2_normalize_many", G2)
fq_from_canonical = _unary("orc_fq_from_canonical", FQ)
fq_to_canonical = _unary("orc_fq_to_canonical", FQ)
fr_from_canonical = _unary("orc_fr_from_canonical", FR)
fr_to_canonical = _unary("orc_fr_to_canonical", FR)


def _with_some(name, unit):
    def f(a, threads=None):
        a, n = _in(a, unit)
        out = np.empty(n * unit, dtype=np.uint8)
        some = np.empty(n, dtype=np.uint8)
        args = [_p(a), _p(out), _p(some), ctypes.c_size_t(n)]
        if threads is not None:
            args.append(ctypes.c_int(threads))
        getattr(lib(), name)(*args)
        return out, some
    return f


fq_inv = _with_some("orc_fq_inv_many", FQ)
fq12_inverse = _with_some("orc_fq12_inverse_many", FQ12)


def final_exp(a, threads=0):
    return _with_some("orc_final_exp_many", FQ12)(a, threads)


def fq12_frobenius(a, power):
    a, n = _in(a, FQ12)
    out = np.empty(n * FQ12, dtype=np.uint8)
    lib().orc_fq12_frobenius_many(_p(a), ctypes.c_int(power), _p(out), ctypes.c_size_t(n))
    return out


def fq12_mul_by_024(f, ell):
    f, n = _in(f, FQ12)
    ell, ne = _in(ell, ELL)
    assert n == ne
    out = np.empty(n * FQ12, dtype=np.uint8)
    lib().orc_fq12_mul_by_024_many(_p(f), _p(ell), _p(out), ctypes.c_size_t(n))
    return out


def _scalar_mul(name, unit):
    def f(p, fr, threads=0):
        p, n = _in(p, unit)
        fr, nf = _in(fr, FR)
        assert n == nf
        out = np.empty(n * unit, dtype=np.uint8)
        getattr(lib(), name)(_p(p), _p(fr), _p(out), ctypes.c_size_t(n), ctypes.c_int(threads))
        return out
    return f


g1_mul = _scalar_mul("orc_g1_mul_many", G1)
g2_mul = _scalar_mul("orc_g2_mul_many", G2)
gt_pow = _scalar_mul("orc_gt_pow_many", FQ12)


def g2_precompute(q):
    q, n = _in(q, G2)
    assert n == 1
    out = np.zeros(G2PRECOMP, dtype=np.uint8)
    lib().orc_g2_precompute.restype = ctypes.c_int
    ok = lib().orc_g2_precompute(_p(q), _p(out))
    return out if ok else None


def pairing_many(p, q, threads=0):
    p, n = _in(p, G1)
    q, nq = _in(q, G2)
    assert n == nq
    out = np.empty(n * FQ12, dtype=np.uint8)
    lib().orc_pairing_many(_p(p), _p(q), _p(out), ctypes.c_size_t(n), ctypes.c_int(threads))
    return out


def pairing_product_many(p, q, pairs_per_check, threads=0):
    """n_checks x pairing_batch(pairs_per_check pairs) -> (Gt images, is_one bytes)."""
    p, n = _in(p, G1)
    q, nq = _in(q, G2)
    assert n == nq
    n_checks = 0 if pairs_per_check == 0 else n // pairs_per_check
    assert pairs_per_check == 0 or n % pairs_per_check == 0
    out = np.empty(n_checks * FQ12, dtype=np.uint8)
    is_one = np.empty(n_checks, dtype=np.uint8)
    lib().orc_pairing_product_many.restype = ctypes.c_int
    rc = lib().orc_pairing_product_many(_p(p), _p(q), ctypes.c_size_t(pairs_per_check), ctypes.c_size_t(n_checks),
                                        _p(out), _p(is_one), ctypes.c_int(threads))
    assert rc == 0
    return out, is_one


def pairing_batch(p, q):
    """One pairing_batch over all pairs (empty -> one)."""
    p, n = _in(p, G1)
    q, nq = _in(q, G2)
    assert n == nq
    out = np.empty(FQ12, dtype=np.uint8)
    is_one = np.empty(1, dtype=np.uint8)
    lib().orc_pairing_product_many(_p(p), _p(q), ctypes.c_size_t(n), ctypes.c_size_t(1), _p(out), _p(is_one), ctypes.c_int(1))
    return out


def miller_loop_batch(q, p):
    """lib::miller_loop_batch: (G2, G1) order.  Returns the Fq12 image or None (Err(ToAffineConversion))."""
    q, n = _in(q, G2)
    p, np_ = _in(p, G1)
    assert n == np_
    out = np.zeros(FQ12, dtype=np.uint8)
    lib().orc_miller_loop_batch.restype = ctypes.c_int
    ok = lib().orc_miller_loop_batch(_p(q), _p(p), ctypes.c_size_t(n), _p(out))
    return out if ok else None


def miller_loop_many(q, p, threads=0):
    q, n = _in(q, G2)
    p, np_ = _in(p, G1)
    assert n == np_
    out = np.empty(n * FQ12, dtype=np.uint8)
    ok = np.empty(n, dtype=np.uint8)
    lib().orc_miller_loop_many(_p(q), _p(p), _p(out), _p(ok), ctypes.c_size_t(n), ctypes.c_int(threads))
    return out, ok


def g1_affine_check(xy):
    """AffineG1::new status per (x, y): 0 Ok, 1 NotOnCurve (src/groups/mod.rs:94-113; no order check for G1)."""
    xy, n = _in(xy, 2 * FQ)
    st = np.empty(n, dtype=np.uint8)
    lib().orc_g1_affine_check_many(_p(xy), _p(st), ctypes.c_size_t(n))
    return st


def g2_affine_check(xy, threads=0):
    """AffineG2::new status per (x, y): 0 Ok, 1 NotOnCurve, 2 NotInSubgroup (src/groups/mod.rs:94-113)."""
    xy, n = _in(xy, 4 * FQ)
    st = np.empty(n, dtype=np.uint8)
    lib().orc_g2_affine_check_many(_p(xy), _p(st), ctypes.c_size_t(n), ctypes.c_int(threads))
    return st


def _raw(a, unit):
    a = np.ascontiguousarray(np.frombuffer(bytes(a), dtype=np.uint8) if isinstance(a, (bytes, bytearray)) else a, dtype=np.uint8).reshape(-1)
    assert a.size % unit == 0
    return a, a.size // unit


def g1_from_compressed(enc):
    """G1::from_compressed per 33-byte record (src/lib.rs:359-375) -> (G1 images, status)."""
    enc, n = _raw(enc, 33)
    out, st = np.empty(n * G1, dtype=np.uint8), np.empty(n, dtype=np.uint8)
    lib().orc_g1_from_compressed_many(_p(enc), _p(out), _p(st), ctypes.c_size_t(n))
    return out, st


def g2_from_compressed(enc, threads=0):
    """G2::from_compressed per 65-byte record (src/lib.rs:506-526) -> (G2 images, status)."""
    enc, n = _raw(enc, 65)
    out, st = np.empty(n * G2, dtype=np.uint8), np.empty(n, dtype=np.uint8)
    lib().orc_g2_from_compressed_many(_p(enc), _p(out), _p(st), ctypes.c_size_t(n), ctypes.c_int(threads))
    return out, st


def u512_divrem_q(value):
    """U512::divrem by the Fq modulus (src/arith.rs:116-138) -> (quotient or None, remainder) as Python ints."""
    be = np.frombuffer(int(value).to_bytes(64, "big"), dtype=np.uint8).copy()
    quot, rem = np.zeros(32, dtype=np.uint8), np.zeros(32, dtype=np.uint8)
    lib().orc_u512_divrem_q.restype = ctypes.c_int
    ok = lib().orc_u512_divrem_q(_p(be), _p(quot), _p(rem))
    return (int.from_bytes(quot.tobytes(), "little") if ok else None), int.from_bytes(rem.tobytes(), "little")


fq_sqrt = _with_some("orc_fq_sqrt_many", FQ)
fq2_sqrt = _with_some("orc_fq2_sqrt_many", 2 * FQ)


def miller_loop_batch_skip(q, p):
    """Un-exponentiated Miller product with pairing_batch's skip semantics (pairs with a point at infinity
    contribute one; empty -> one): product of the single Miller values, src/groups/mod.rs:609-640."""
    vals, _ok = miller_loop_many(q, p)
    acc = np.frombuffer(((1 << 256) % 0x30644e72e131a029b85045b68181585d97816a916871ca8d3c208c16d87cfd47).to_bytes(32, "little") + bytes(352),
                        dtype=np.uint8).copy()
    for i in range(vals.size // FQ12):
        acc = fq12_mul(acc, vals[i * FQ12:(i + 1) * FQ12])
    return acc


def constants():
    """Raw dump of the derived constants (order: see orc_constants in bn254_oracle.c)."""
    buf = np.zeros(4096, dtype=np.uint8)
    lib().orc_constants(_p(buf))
    return buf
