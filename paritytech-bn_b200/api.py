"""Host-side mirror of the substrate-bn pairing surface over the bn254_b200 C ABI (ctypes).

This module is the Python test/bench host for libbn254_b200.so (include/bn254_b200.h).  The names
follow the crate's public API (src/lib.rs): pairing (:611), pairing_batch (:615), miller_loop_batch
(:625, note the (G2, G1) order), Gt.final_exponentiation (:598), Gt == / * / pow / inverse
(:584-609), G1 * Fr (:425), G2 * Fr (:576), Group::normalize (:391, :542) -- all batched.

Values are numpy uint8 arrays holding the crate's #[repr(C)] Montgomery images (Fr/Fq 32 B,
G1 96 B, G2 192 B, Gt 384 B).  There is NO CPU fallback: if the CUDA library is missing or no
GPU is present, construction of an Engine raises.
"""
import ctypes
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))

FR_BYTES, FQ_BYTES, G1_BYTES, G2_BYTES, GT_BYTES = 32, 32, 96, 192, 384
G2PRECOMP_BYTES = 128 + 87 * 192
ERR_TO_AFFINE = 1

FIELD_OPS = {"fq_mul": (0, 32), "fq_add": (1, 32), "fq_sub": (2, 32), "fq2_mul": (3, 64), "fq2_sqr": (4, 64),
             "fq6_mul": (5, 192), "fq12_mul": (6, 384), "fq12_sqr": (7, 384), "fq12_cyclotomic_sqr": (8, 384),
             "fq_inv": (9, 32), "fq2_mul_by_xi": (10, 64),
             # the slot-file bodies of csrc/bn_slots.cuh, one call at a time
             "slot_dot3": (11, 192), "slot_fq12_mul": (12, 384), "slot_fq12_sqr": (13, 384),
             "slot_cyclotomic_sqr": (14, 384), "slot_fq6_mul": (15, 192)}


class BnError(RuntimeError):
    pass


class CurveError(BnError):
    """CurveError::ToAffineConversion (src/lib.rs:109-115)."""


def library_path(variant=None):
    variant = variant if variant is not None else os.environ.get("BN254_B200_VARIANT", "")
    return os.path.join(_HERE, "lib", "libbn254_b200%s.so" % variant)


_libs = {}


def load_library(variant=None):
    path = library_path(variant)
    if path not in _libs:
        if not os.path.exists(path):
            raise BnError("CUDA library not built: %s (run `python -c 'import __graft_entry__ as g; g.build()'`); "
                          "there is no CPU fallback" % path)
        lib = ctypes.CDLL(path)
        lib.bn254_last_error.restype = ctypes.c_char_p
        lib.bn254_build_info.restype = ctypes.c_char_p
        lib.bn254_ctx_stream.restype = ctypes.c_void_p
        lib.bn254_ctx_launch_count.restype = ctypes.c_uint64
        lib.bn254_g2precomp_count.restype = ctypes.c_size_t
        lib.bn254_ctx_destroy.restype = None
        lib.bn254_g2precomp_destroy.restype = None
        _libs[path] = lib
    return _libs[path]


def _arr(a, unit):
    """Byte view of a buffer of `unit`-byte elements.  Arrays of wider integers (the natural uint32/uint64 limb layout of a
    U256) are REINTERPRETED as their little-endian bytes, never value-cast; anything that is not an integer array is refused."""
    if isinstance(a, (bytes, bytearray, memoryview)):
        a = np.frombuffer(bytes(a), dtype=np.uint8)
    a = np.asarray(a)
    if a.dtype != np.uint8:
        if a.dtype.kind not in "ui":
            raise TypeError("expected a byte buffer or an integer limb array, got dtype %s" % a.dtype)
        if a.dtype.byteorder == ">":
            raise TypeError("big-endian limb arrays are not the crate's memory image")
        a = np.ascontiguousarray(a).view(np.uint8)
    a = np.ascontiguousarray(a).reshape(-1)
    if a.size % unit:
        raise ValueError("buffer of %d bytes is not a whole number of %d-byte elements" % (a.size, unit))
    return a, a.size // unit


def _p(a):
    return a.ctypes.data_as(ctypes.c_void_p) if a is not None else None


def _out(out, nbytes):
    """Caller-owned result buffer (e.g. pinned memory) or a fresh array."""
    if out is None:
        return np.empty(nbytes, dtype=np.uint8)
    assert out.dtype == np.uint8 and out.size == nbytes and out.flags["C_CONTIGUOUS"]
    return out


class CtxOptions(ctypes.Structure):
    """bn254_ctx_options (include/bn254_b200.h): -1 = library default."""
    _fields_ = [("struct_size", ctypes.c_uint32)] + [(k, ctypes.c_int32) for k in (
        "pipeline", "big_blocks", "fused_miller", "split_fexp", "fexp_smem", "miller_smem", "lines_smem", "subgroup_ref",
        "zerocopy_out", "coop_max_n", "stack_bytes", "mul_compact", "coop4_max_n")]

    def __init__(self, **kw):
        super().__init__()
        self.struct_size = ctypes.sizeof(CtxOptions)
        for k, _ in self._fields_[1:]:
            setattr(self, k, -1)
        for k, v in kw.items():
            if k not in dict(self._fields_):
                raise KeyError("unknown context option %r" % k)
            setattr(self, k, int(v))


class G2Precomp:
    """Device-resident G2Precomp tables (engine-side handle; src/groups/mod.rs:573-577)."""

    def __init__(self, engine, handle):
        self.engine, self.handle = engine, handle

    def __len__(self):
        return self.engine.lib.bn254_g2precomp_count(self.handle)

    def export(self, i):
        out = np.empty(G2PRECOMP_BYTES, dtype=np.uint8)
        self.engine._check(self.engine.lib.bn254_g2precomp_export(self.engine.ctx, self.handle, ctypes.c_size_t(i), _p(out)))
        return out

    def close(self):
        if self.handle:
            self.engine.lib.bn254_g2precomp_destroy(self.handle)
            self.handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class Engine:
    """One CUDA device + stream (bn254_ctx)."""

    def __init__(self, device=0, variant=None, **options):
        """options: fields of bn254_ctx_options (e.g. fexp_smem=0, big_blocks=2, coop_max_n=0); unset fields keep the
        library defaults (which the BN254_* environment variables may override)."""
        self.lib = load_library(variant)
        self.ctx = ctypes.c_void_p()
        if options:
            opt = CtxOptions(**options)
            rc = self.lib.bn254_ctx_create_ex(ctypes.c_int(device), ctypes.byref(opt), ctypes.byref(self.ctx))
        else:
            rc = self.lib.bn254_ctx_create(ctypes.c_int(device), ctypes.byref(self.ctx))
        if rc != 0:
            raise BnError("bn254_ctx_create(%d) failed: %s" % (device, self.lib.bn254_last_error().decode()))
        self.device = device

    def close(self):
        if self.ctx:
            self.lib.bn254_ctx_destroy(self.ctx)
            self.ctx = ctypes.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _check(self, rc):
        if rc == ERR_TO_AFFINE:
            raise CurveError("ToAffineConversion")
        if rc != 0:
            raise BnError("bn254 error %d: %s" % (rc, self.lib.bn254_last_error().decode()))

    @property
    def stream(self):
        return self.lib.bn254_ctx_stream(self.ctx)

    def launch_count(self):
        return int(self.lib.bn254_ctx_launch_count(self.ctx))

    def synchronize(self):
        self._check(self.lib.bn254_ctx_synchronize(self.ctx))

    def build_info(self):
        return self.lib.bn254_build_info().decode()

    # ---- pairing surface (host buffers) ----------------------------------------------------------
    def pairing(self, g1, g2, out=None):
        """n x pairing(p, q) -> n x Gt (src/lib.rs:611-613).  `out`: optional caller-owned uint8 buffer of
        n*384 bytes (e.g. pinned memory) to receive the result."""
        g1, n = _arr(g1, G1_BYTES)
        g2, n2 = _arr(g2, G2_BYTES)
        assert n == n2
        if out is None:
            out = np.empty(n * GT_BYTES, dtype=np.uint8)
        else:
            assert out.dtype == np.uint8 and out.size == n * GT_BYTES and out.flags["C_CONTIGUOUS"]
        self._check(self.lib.bn254_pairing_many(self.ctx, _p(g1), _p(g2), _p(out), ctypes.c_size_t(n)))
        return out

    def pairing_batch(self, g1, g2, pairs_per_check=None, out=None, is_one=None):
        """pairing_batch (src/lib.rs:615-623).  With pairs_per_check: many independent checks.
        Returns (Gt images, is_one bytes)."""
        g1, n = _arr(g1, G1_BYTES)
        g2, n2 = _arr(g2, G2_BYTES)
        assert n == n2
        if pairs_per_check is None:
            pairs_per_check, n_checks = n, 1
        else:
            assert pairs_per_check > 0 and n % pairs_per_check == 0
            n_checks = n // pairs_per_check
        out = _out(out, n_checks * GT_BYTES)
        is_one = _out(is_one, n_checks)
        self._check(self.lib.bn254_pairing_product_many(self.ctx, _p(g1), _p(g2), ctypes.c_size_t(pairs_per_check),
                                                        ctypes.c_size_t(n_checks), _p(out), _p(is_one)))
        return out, is_one

    def miller_loop_batch(self, g2, g1):
        """miller_loop_batch(&[(G2, G1)]) (src/lib.rs:625-633); raises CurveError on a point at infinity."""
        g2, n = _arr(g2, G2_BYTES)
        g1, n1 = _arr(g1, G1_BYTES)
        assert n == n1
        out = np.empty(GT_BYTES, dtype=np.uint8)
        self._check(self.lib.bn254_miller_loop_batch(self.ctx, _p(g2), _p(g1), ctypes.c_size_t(n), _p(out)))
        return out

    def miller_product(self, g1, g2):
        """Un-exponentiated product of the Miller values of all pairs, pairing_batch skip semantics
        (the per-GPU partial of a sharded product)."""
        g1, n = _arr(g1, G1_BYTES)
        g2, n2 = _arr(g2, G2_BYTES)
        assert n == n2
        out = np.empty(GT_BYTES, dtype=np.uint8)
        self._check(self.lib.bn254_miller_product(self.ctx, _p(g1) if n else None, _p(g2) if n else None, ctypes.c_size_t(n), _p(out)))
        return out

    def gt_one(self):
        """Gt::one() (src/lib.rs:588) = the empty product."""
        e = np.empty(0, dtype=np.uint8)
        return self.miller_product(e, e)

    def miller_loop_many(self, g2, g1):
        g2, n = _arr(g2, G2_BYTES)
        g1, n1 = _arr(g1, G1_BYTES)
        assert n == n1
        out = np.empty(n * GT_BYTES, dtype=np.uint8)
        ok = np.empty(n, dtype=np.uint8)
        self._check(self.lib.bn254_miller_loop_many(self.ctx, _p(g2), _p(g1), _p(out), _p(ok), ctypes.c_size_t(n)))
        return out, ok

    def final_exponentiation(self, gt):
        gt, n = _arr(gt, GT_BYTES)
        out = np.empty(n * GT_BYTES, dtype=np.uint8)
        some = np.empty(n, dtype=np.uint8)
        self._check(self.lib.bn254_final_exp_many(self.ctx, _p(gt), _p(out), _p(some), ctypes.c_size_t(n)))
        return out, some

    def gt_mul(self, a, b):
        a, n = _arr(a, GT_BYTES)
        b, n2 = _arr(b, GT_BYTES)
        assert n == n2
        out = np.empty(n * GT_BYTES, dtype=np.uint8)
        self._check(self.lib.bn254_gt_mul_many(self.ctx, _p(a), _p(b), _p(out), ctypes.c_size_t(n)))
        return out

    def gt_pow(self, a, fr):
        a, n = _arr(a, GT_BYTES)
        fr, n2 = _arr(fr, FR_BYTES)
        assert n == n2
        out = np.empty(n * GT_BYTES, dtype=np.uint8)
        self._check(self.lib.bn254_gt_pow_many(self.ctx, _p(a), _p(fr), _p(out), ctypes.c_size_t(n)))
        return out

    def gt_inverse(self, a):
        a, n = _arr(a, GT_BYTES)
        out = np.empty(n * GT_BYTES, dtype=np.uint8)
        some = np.empty(n, dtype=np.uint8)
        self._check(self.lib.bn254_gt_inverse_many(self.ctx, _p(a), _p(out), _p(some), ctypes.c_size_t(n)))
        return out, some

    def gt_eq(self, a, b):
        a, n = _arr(a, GT_BYTES)
        b, n2 = _arr(b, GT_BYTES)
        assert n == n2
        eq = np.empty(n, dtype=np.uint8)
        self._check(self.lib.bn254_gt_eq_many(self.ctx, _p(a), _p(b), _p(eq), ctypes.c_size_t(n)))
        return eq

    def g1_mul(self, g1, fr, out=None):
        g1, n = _arr(g1, G1_BYTES)
        fr, n2 = _arr(fr, FR_BYTES)
        assert n == n2
        out = _out(out, n * G1_BYTES)
        self._check(self.lib.bn254_g1_mul_many(self.ctx, _p(g1), _p(fr), _p(out), ctypes.c_size_t(n)))
        return out

    def g2_mul(self, g2, fr, out=None):
        g2, n = _arr(g2, G2_BYTES)
        fr, n2 = _arr(fr, FR_BYTES)
        assert n == n2
        out = _out(out, n * G2_BYTES)
        self._check(self.lib.bn254_g2_mul_many(self.ctx, _p(g2), _p(fr), _p(out), ctypes.c_size_t(n)))
        return out

    def g1_mul_normalized(self, g1, fr, out=None):
        """(p * k).normalize() in one call (signed-window kernel; bytes of the reference's normalised result)."""
        g1, n = _arr(g1, G1_BYTES)
        fr, n2 = _arr(fr, FR_BYTES)
        assert n == n2
        out = _out(out, n * G1_BYTES)
        self._check(self.lib.bn254_g1_mul_normalized_many(self.ctx, _p(g1), _p(fr), _p(out), ctypes.c_size_t(n)))
        return out

    def g2_mul_normalized(self, g2, fr, out=None):
        g2, n = _arr(g2, G2_BYTES)
        fr, n2 = _arr(fr, FR_BYTES)
        assert n == n2
        out = _out(out, n * G2_BYTES)
        self._check(self.lib.bn254_g2_mul_normalized_many(self.ctx, _p(g2), _p(fr), _p(out), ctypes.c_size_t(n)))
        return out

    def _group_op(self, which, op, a, b=None):
        unit = G1_BYTES if which == "g1" else G2_BYTES
        a, n = _arr(a, unit)
        fn = getattr(self.lib, "bn254_%s_%s_many" % (which, op))
        if op in ("add", "eq"):
            b, n2 = _arr(b, unit)
            assert n == n2
            out = np.empty(n if op == "eq" else n * unit, dtype=np.uint8)
            self._check(fn(self.ctx, _p(a), _p(b), _p(out), ctypes.c_size_t(n)))
        else:
            out = np.empty(n * unit, dtype=np.uint8)
            self._check(fn(self.ctx, _p(a), _p(out), ctypes.c_size_t(n)))
        return out

    def g1_add(self, a, b):
        """G1 + G1 elementwise (src/lib.rs:405-411): the reference's Jacobian representative."""
        return self._group_op("g1", "add", a, b)

    def g2_add(self, a, b):
        return self._group_op("g2", "add", a, b)

    def g1_double(self, a):
        return self._group_op("g1", "double", a)

    def g2_double(self, a):
        return self._group_op("g2", "double", a)

    def g1_neg(self, a):
        return self._group_op("g1", "neg", a)

    def g2_neg(self, a):
        return self._group_op("g2", "neg", a)

    def g1_eq(self, a, b):
        """G1 == G1 elementwise: projective equality (src/groups/mod.rs:169-195), one byte per element."""
        return self._group_op("g1", "eq", a, b)

    def g2_eq(self, a, b):
        return self._group_op("g2", "eq", a, b)

    def g1_normalize(self, g1):
        g1, n = _arr(g1, G1_BYTES)
        out = np.empty(n * G1_BYTES, dtype=np.uint8)
        self._check(self.lib.bn254_g1_normalize_many(self.ctx, _p(g1), _p(out), ctypes.c_size_t(n)))
        return out

    def g2_normalize(self, g2):
        g2, n = _arr(g2, G2_BYTES)
        out = np.empty(n * G2_BYTES, dtype=np.uint8)
        self._check(self.lib.bn254_g2_normalize_many(self.ctx, _p(g2), _p(out), ctypes.c_size_t(n)))
        return out

    def g1_affine_check(self, xy):
        """AffineG1::new per (x, y) pair of Fq images: 0 Ok, 1 NotOnCurve (src/groups/mod.rs:94-113)."""
        xy, n = _arr(xy, 2 * FQ_BYTES)
        st = np.empty(n, dtype=np.uint8)
        self._check(self.lib.bn254_g1_affine_check_many(self.ctx, _p(xy), _p(st), ctypes.c_size_t(n)))
        return st

    def g2_affine_check(self, xy):
        """AffineG2::new per (x, y) pair of Fq2 images: 0 Ok, 1 NotOnCurve, 2 NotInSubgroup."""
        xy, n = _arr(xy, 4 * FQ_BYTES)
        st = np.empty(n, dtype=np.uint8)
        self._check(self.lib.bn254_g2_affine_check_many(self.ctx, _p(xy), _p(st), ctypes.c_size_t(n)))
        return st

    def _from_compressed(self, fn, enc, rec, unit):
        enc = np.ascontiguousarray(np.frombuffer(bytes(enc), dtype=np.uint8) if isinstance(enc, (bytes, bytearray)) else enc,
                                   dtype=np.uint8).reshape(-1)
        if enc.size % rec:
            raise ValueError("encoding length is not a multiple of %d bytes" % rec)
        n = enc.size // rec
        out, st = np.empty(n * unit, dtype=np.uint8), np.empty(n, dtype=np.uint8)
        self._check(fn(self.ctx, _p(enc), _p(out), _p(st), ctypes.c_size_t(n)))
        return out, st

    def g1_from_compressed(self, enc):
        """G1::from_compressed per 33-byte record (src/lib.rs:359-375) -> (G1 images, status bytes)."""
        return self._from_compressed(self.lib.bn254_g1_from_compressed_many, enc, 33, G1_BYTES)

    def g2_from_compressed(self, enc):
        """G2::from_compressed per 65-byte record (src/lib.rs:506-526) -> (G2 images, status bytes)."""
        return self._from_compressed(self.lib.bn254_g2_from_compressed_many, enc, 65, G2_BYTES)

    # ---- shared G2Precomp -------------------------------------------------------------------------
    def g2_precompute(self, g2):
        g2, n = _arr(g2, G2_BYTES)
        h = ctypes.c_void_p()
        self._check(self.lib.bn254_g2_precompute(self.ctx, _p(g2), ctypes.c_size_t(n), ctypes.byref(h)))
        return G2Precomp(self, h)

    def pairing_batch_precomp(self, precomp, g1_shared, g1_var, g2_var, v, out=None, is_one=None):
        """Groth16-shaped checks: per check len(precomp) shared-table pairs + v variable pairs."""
        m = len(precomp) if precomp is not None else 0
        g1_var, nv = _arr(g1_var, G1_BYTES)
        g2_var, nv2 = _arr(g2_var, G2_BYTES)
        g1_shared, ns = _arr(g1_shared, G1_BYTES)
        assert nv == nv2
        n_checks = ns // m if m else nv // v
        assert ns == n_checks * m and nv == n_checks * v
        out = _out(out, n_checks * GT_BYTES)
        is_one = _out(is_one, n_checks)
        self._check(self.lib.bn254_pairing_product_precomp_many(
            self.ctx, precomp.handle if precomp is not None else None, _p(g1_shared), _p(g1_var), _p(g2_var),
            ctypes.c_size_t(v), ctypes.c_size_t(n_checks), _p(out), _p(is_one)))
        return out, is_one

    # ---- device-pointer entry points (inputs already resident in HBM; asynchronous) ----------------
    def pairing_dev(self, d_g1, d_g2, d_out, n):
        self._check(self.lib.bn254_pairing_many_dev(self.ctx, ctypes.c_void_p(d_g1), ctypes.c_void_p(d_g2),
                                                    ctypes.c_void_p(d_out), ctypes.c_size_t(n)))

    def pairing_batch_dev(self, d_g1, d_g2, pairs_per_check, n_checks, d_out, d_is_one):
        self._check(self.lib.bn254_pairing_product_many_dev(self.ctx, ctypes.c_void_p(d_g1), ctypes.c_void_p(d_g2),
                                                            ctypes.c_size_t(pairs_per_check), ctypes.c_size_t(n_checks),
                                                            ctypes.c_void_p(d_out), ctypes.c_void_p(d_is_one)))

    def pairing_batch_precomp_dev(self, precomp, d_g1s, d_g1v, d_g2v, v, n_checks, d_out, d_is_one):
        self._check(self.lib.bn254_pairing_product_precomp_many_dev(
            self.ctx, precomp.handle, ctypes.c_void_p(d_g1s), ctypes.c_void_p(d_g1v), ctypes.c_void_p(d_g2v),
            ctypes.c_size_t(v), ctypes.c_size_t(n_checks), ctypes.c_void_p(d_out), ctypes.c_void_p(d_is_one)))

    def g1_mul_dev(self, d_g1, d_fr, d_out, n):
        self._check(self.lib.bn254_g1_mul_many_dev(self.ctx, ctypes.c_void_p(d_g1), ctypes.c_void_p(d_fr),
                                                   ctypes.c_void_p(d_out), ctypes.c_size_t(n)))

    def g2_mul_dev(self, d_g2, d_fr, d_out, n):
        self._check(self.lib.bn254_g2_mul_many_dev(self.ctx, ctypes.c_void_p(d_g2), ctypes.c_void_p(d_fr),
                                                   ctypes.c_void_p(d_out), ctypes.c_size_t(n)))

    def g1_mul_normalized_dev(self, d_g1, d_fr, d_out, n):
        self._check(self.lib.bn254_g1_mul_normalized_many_dev(self.ctx, ctypes.c_void_p(d_g1), ctypes.c_void_p(d_fr),
                                                              ctypes.c_void_p(d_out), ctypes.c_size_t(n)))

    def g2_mul_normalized_dev(self, d_g2, d_fr, d_out, n):
        self._check(self.lib.bn254_g2_mul_normalized_many_dev(self.ctx, ctypes.c_void_p(d_g2), ctypes.c_void_p(d_fr),
                                                              ctypes.c_void_p(d_out), ctypes.c_size_t(n)))

    def miller_product_dev(self, d_g1, d_g2, n, d_out):
        """Un-exponentiated product of n device-resident pairs into d_out (384 B); asynchronous."""
        self._check(self.lib.bn254_miller_product_dev(self.ctx, ctypes.c_void_p(d_g1), ctypes.c_void_p(d_g2), ctypes.c_size_t(n),
                                                      ctypes.c_void_p(d_out)))

    def pairing_timed_dev(self, d_g1, d_g2, d_out, n):
        """bn254_pairing_many_dev_timed: per-kernel milliseconds [affine, lines, miller, easy, hard]."""
        ms = (ctypes.c_float * 5)()
        self._check(self.lib.bn254_pairing_many_dev_timed(self.ctx, ctypes.c_void_p(d_g1), ctypes.c_void_p(d_g2), ctypes.c_void_p(d_out),
                                                          ctypes.c_size_t(n), ms))
        return [float(x) for x in ms]

    def g1_from_compressed_dev(self, d_enc, d_out, d_status, n):
        self._check(self.lib.bn254_g1_from_compressed_many_dev(self.ctx, ctypes.c_void_p(d_enc), ctypes.c_void_p(d_out),
                                                               ctypes.c_void_p(d_status), ctypes.c_size_t(n)))

    def g2_from_compressed_dev(self, d_enc, d_out, d_status, n):
        self._check(self.lib.bn254_g2_from_compressed_many_dev(self.ctx, ctypes.c_void_p(d_enc), ctypes.c_void_p(d_out),
                                                               ctypes.c_void_p(d_status), ctypes.c_size_t(n)))

    def g1_affine_check_dev(self, d_xy, d_status, n):
        self._check(self.lib.bn254_g1_affine_check_many_dev(self.ctx, ctypes.c_void_p(d_xy), ctypes.c_void_p(d_status), ctypes.c_size_t(n)))

    def g2_affine_check_dev(self, d_xy, d_status, n):
        self._check(self.lib.bn254_g2_affine_check_many_dev(self.ctx, ctypes.c_void_p(d_xy), ctypes.c_void_p(d_status), ctypes.c_size_t(n)))

    # ---- test hooks / microbenchmarks ------------------------------------------------------------------
    def field_op(self, name, a, b=None):
        op, unit = FIELD_OPS[name]
        a, n = _arr(a, unit)
        if b is not None:
            b, n2 = _arr(b, unit)
            assert n == n2
        out = np.empty(n * unit, dtype=np.uint8)
        self._check(self.lib.bn254_field_op_many(self.ctx, ctypes.c_int(op), _p(a), _p(b), _p(out), ctypes.c_size_t(n)))
        return out

    def microbench(self, kind, grid, block, iters):
        ms, ops = ctypes.c_double(), ctypes.c_double()
        self._check(self.lib.bn254_microbench(self.ctx, ctypes.c_int(kind), ctypes.c_int(grid), ctypes.c_int(block),
                                              ctypes.c_int(iters), ctypes.byref(ms), ctypes.byref(ops)))
        return ms.value, ops.value


def _ctx_list(engines):
    return (ctypes.c_void_p * len(engines))(*[e.ctx.value if isinstance(e.ctx, ctypes.c_void_p) else e.ctx for e in engines])


def pairing_product_sharded(engines, g1, g2, timed=False):
    """One product over all pairs, sharded by index range over several engines (one per GPU);
    384-byte partials gathered to engines[0] by peer copy, one final exponentiation.
    timed=True also returns [miller_ms, gather_ms, combine_ms] (host clock, extra synchronisations)."""
    g1, n = _arr(g1, G1_BYTES)
    g2, n2 = _arr(g2, G2_BYTES)
    assert n == n2 and engines
    lib = engines[0].lib
    ctxs = _ctx_list(engines)
    out = np.empty(GT_BYTES, dtype=np.uint8)
    is_one = np.empty(1, dtype=np.uint8)
    if timed:
        ms = (ctypes.c_double * 3)()
        engines[0]._check(lib.bn254_pairing_product_sharded_timed(ctxs, ctypes.c_size_t(len(engines)), _p(g1), _p(g2),
                                                                  ctypes.c_size_t(n), _p(out), _p(is_one), ms))
        return out, bool(is_one[0]), [float(x) for x in ms]
    engines[0]._check(lib.bn254_pairing_product_sharded(ctxs, ctypes.c_size_t(len(engines)), _p(g1), _p(g2),
                                                        ctypes.c_size_t(n), _p(out), _p(is_one)))
    return out, bool(is_one[0])


def pairing_sharded(engines, g1, g2, out=None):
    """n independent pairings in ONE call over a device list (bn254_pairing_many_sharded): contiguous index ranges,
    one host thread per engine, results written to the caller's index ranges."""
    g1, n = _arr(g1, G1_BYTES)
    g2, n2 = _arr(g2, G2_BYTES)
    assert n == n2 and engines
    if out is None:
        out = np.empty(n * GT_BYTES, dtype=np.uint8)
    else:
        assert out.dtype == np.uint8 and out.size == n * GT_BYTES and out.flags["C_CONTIGUOUS"]
    engines[0]._check(engines[0].lib.bn254_pairing_many_sharded(_ctx_list(engines), ctypes.c_size_t(len(engines)), _p(g1), _p(g2),
                                                                _p(out), ctypes.c_size_t(n)))
    return out
