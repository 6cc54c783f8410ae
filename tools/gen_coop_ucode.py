#!/usr/bin/env python3
"""Microcode of the warp-cooperative BN254 pairing (latency path, csrc/bn_coop.cuh) -- generator, simulator, emitter.

One WARP computes one pairing.  The warp is a 32-lane vector machine over Fq2: a per-warp file of Fq2 slots in shared memory
and two instruction kinds, executed in lock step by all lanes, each lane with its own operands:

  MUL  dst = conjA?(sum_{t<na} +-slot[a_t])  *  conjB?(sum_{t<nb} +-slot[b_t])                  (one Fq2 product per lane)
  LIN  dst = half?( P + xi * X ),  P = sum of the terms of group 0, X = sum of the terms of group 1 (xi = 9 + u)
  INV  dst = (slot[a].c0)^-1 as (Fq, 0)   (one lane, Fermat in registers; everything else is MUL/LIN)

At most 8 terms per lane and step (na + nb, or |P| + |X|); small multiples are repeated terms.  A routine is a short list of steps; a program is a list of calls (routine, base D, base A, base B): operand
addresses are absolute slots (sel 0) or relative to one of the three bases of the call (sel 1..3), so one MUL12 routine serves
every Fq12 product of the final exponentiation.  The CUDA side is a generic executor with no knowledge of the algorithm;
everything algorithmic lives here, where the simulator below runs the very same tables on Python integers against the oracle
(tests/test_coop_ucode.py), so the schedule is proven on the CPU and the GPU run only has to confirm the executor.

Formulas follow the reference exactly as the throughput kernels do: G2 line steps src/groups/mod.rs:731-776, precompute /
Miller loop :580-606, :701-728, Fq12 tower src/fields/fq12.rs (mul :316-328, squared :295-303, mul_by_024 :130-196,
cyclotomic_squared :198-247, final exponentiation :62-110, exp_by_neg_z :121-124 with the width-4 signed window of
csrc/bn_tower.cuh), fq6 inverse src/fields/fq6.rs:179-191.

usage: gen_coop_ucode.py            write paritytech-bn_b200/csrc/bn_coop_ucode.cuh
       gen_coop_ucode.py --check    simulate a few pairings against the oracle
"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
from oracle import bn254_ref as o  # noqa: E402

Q = o.Q
OP_MUL, OP_LIN, OP_INV = 1, 2, 3
NT = 8                     # term fields per lane and step (MUL: 4 + 4, LIN: 8)
NO_DST = 0x3ff             # inactive lane


# ------------------------------------------------------------------------------------------------------------------
# operands
# ------------------------------------------------------------------------------------------------------------------
class Ref:
    """A slot address: absolute (sel 0) or relative to base D (1), A (2), B (3) of the call."""
    __slots__ = ("sel", "off")

    def __init__(self, sel, off):
        assert 0 <= sel < 4 and 0 <= off < 256
        self.sel, self.off = sel, off

    def code(self):
        return (self.sel << 8) | self.off

    def __add__(self, k):
        return Ref(self.sel, self.off + k)

    def __repr__(self):
        return "%s%d" % ("@DAB"[self.sel], self.off)


def S(off):
    return Ref(0, off)


D, A, B = Ref(1, 0), Ref(2, 0), Ref(3, 0)


def T(ref, sign=1, grp=0):
    """One term: +-slot (group 1 = multiplied by xi in a LIN step)."""
    assert sign in (1, -1) and grp in (0, 1)
    return (ref, sign, grp)


def neg(t):
    return (t[0], -t[1], t[2])


def xi(t):
    return (t[0], t[1], 1)


def times(t, m):
    """m * t as repeated terms."""
    return [t] * m


def terms(x):
    """Normalise: a Ref, a term or a list of either -> list of terms."""
    if isinstance(x, Ref):
        return [T(x)]
    if isinstance(x, tuple) and len(x) == 3 and isinstance(x[0], Ref):
        return [x]
    out = []
    for e in x:
        out.extend(terms(e))
    return out


# ------------------------------------------------------------------------------------------------------------------
# assembler
# ------------------------------------------------------------------------------------------------------------------
class Step:
    def __init__(self, op):
        self.op, self.lanes = op, []      # lanes: dicts

    def counts(self):
        """(n0, n1): the uniform term counts of the step -- MUL: A side / B side, LIN: group 0 / group 1, INV: (1, 0)."""
        if self.op == OP_MUL:
            return max(len(l["a"]) for l in self.lanes), max(len(l["b"]) for l in self.lanes)
        if self.op == OP_LIN:
            return (max(sum(1 for t in l["t"] if t[2] == 0) for l in self.lanes),
                    max(sum(1 for t in l["t"] if t[2] == 1) for l in self.lanes))
        return 1, 0

    def header(self):
        """op | n0 << 4 | n1 << 8 | any conj << 12 | any half << 13"""
        n0, n1 = self.counts()
        assert n0 + n1 <= NT, (n0, n1)
        cj = any(l.get("ca") or l.get("cb") for l in self.lanes)
        hf = any(l.get("half") for l in self.lanes)
        return self.op | (n0 << 4) | (n1 << 8) | (int(cj) << 12) | (int(hf) << 13)


class Routine:
    def __init__(self, name):
        self.name, self.steps = name, []

    def mul(self, lanes):
        """lanes: list of (dst, A-side, B-side[, conjA, conjB])."""
        st = Step(OP_MUL)
        for ln in lanes:
            dst, a, b = ln[0], terms(ln[1]), terms(ln[2])
            ca, cb = (ln[3], ln[4]) if len(ln) > 3 else (False, False)
            assert 1 <= len(a) and 1 <= len(b), (self.name, len(a), len(b))
            assert all(t[2] == 0 for t in a + b), "xi groups exist in LIN steps only"
            st.lanes.append(dict(dst=dst, a=a, b=b, ca=bool(ca), cb=bool(cb)))
        self._add(st)

    def lin(self, lanes):
        """lanes: list of (dst, terms[, half])."""
        st = Step(OP_LIN)
        for ln in lanes:
            dst, t = ln[0], terms(ln[1])
            assert 1 <= len(t) <= NT, (self.name, len(t))
            t = [x for x in t if x[2] == 0] + [x for x in t if x[2] == 1]          # group 0 first
            st.lanes.append(dict(dst=dst, t=t, half=bool(ln[2]) if len(ln) > 2 else False))
        self._add(st)

    def inv(self, dst, src):
        st = Step(OP_INV)
        st.lanes.append(dict(dst=dst, t=terms(src)))
        self._add(st)

    def _add(self, st):
        assert 1 <= len(st.lanes) <= 32, (self.name, len(st.lanes))
        st.header()                                                   # term budget
        # one synchronisation per step: no lane may read a slot that ANOTHER lane of the same step writes
        # (compared as (sel, off): two different bases may alias only where the call list says so -- checked by the simulator)
        for i, li in enumerate(st.lanes):
            srcs = [t[0].code() for k in ("a", "b", "t") for t in li.get(k, [])]
            for j, lj in enumerate(st.lanes):
                if i != j:
                    assert lj["dst"].code() not in srcs, "%s: lane %d reads the destination of lane %d" % (self.name, i, j)
            for j in range(i):
                assert st.lanes[j]["dst"].code() != li["dst"].code(), "%s: two lanes write one slot" % self.name
        self.steps.append(st)


# ------------------------------------------------------------------------------------------------------------------
# simulator (Python integers; the semantics the CUDA executor implements)
# ------------------------------------------------------------------------------------------------------------------
def _resolve(ref, bases):
    return ref.off + (0 if ref.sel == 0 else bases[ref.sel - 1])


def _sum(slots, tl, bases, grp=None):
    acc = (0, 0)
    for ref, sign, g in tl:
        if grp is not None and g != grp:
            continue
        v = slots[_resolve(ref, bases)]
        acc = o.fq2_add(acc, v) if sign > 0 else o.fq2_sub(acc, v)
    return acc


def _conj(v):
    return (v[0], (-v[1]) % Q)


def run_step(slots, st, bases, check_alias=True):
    writes = {}
    reads = {}
    for i, ln in enumerate(st.lanes):
        for k in ("a", "b", "t"):
            for t in ln.get(k, []):
                reads.setdefault(_resolve(t[0], bases), set()).add(i)
        if st.op == OP_MUL:
            a = _sum(slots, ln["a"], bases)
            b = _sum(slots, ln["b"], bases)
            if ln["ca"]:
                a = _conj(a)
            if ln["cb"]:
                b = _conj(b)
            r = o.fq2_mul(a, b)
        elif st.op == OP_LIN:
            p = _sum(slots, ln["t"], bases, 0)
            x = _sum(slots, ln["t"], bases, 1)
            r = o.fq2_add(p, o.fq2_mul_by_nonresidue(x))
            if ln["half"]:
                h = pow(2, -1, Q)
                r = (r[0] * h % Q, r[1] * h % Q)
        else:
            v = _sum(slots, ln["t"], bases)
            r = (pow(v[0], Q - 2, Q), 0)
        d = _resolve(ln["dst"], bases)
        assert d not in writes, "two lanes write slot %d" % d
        writes[d] = (i, r)
    if check_alias:
        for d, (i, _) in writes.items():
            assert reads.get(d, set()) <= {i}, "slot %d written by lane %d is read by lanes %s in the same step" % (d, i, reads[d])
    for d, (_, r) in writes.items():
        slots[d] = r


def run_program(slots, routines, calls):
    for rid, bd, ba, bb in calls:
        for st in routines[rid].steps:
            run_step(slots, st, (bd, ba, bb))


# ------------------------------------------------------------------------------------------------------------------
# slot map (absolute addresses of the per-warp file; 64 bytes per slot)
# ------------------------------------------------------------------------------------------------------------------
ZERO, ONE, C_B3, C_TWQX, C_TWQY = S(0), S(1), S(2), S(3), S(4)
C_F6C1 = {1: S(5), 2: S(6), 3: S(7)}
C_F6C2 = {1: S(8), 2: S(9), 3: S(10)}
C_F12C = {1: S(11), 2: S(12), 3: S(13)}
N_CONST = 14
IN_P, IN_Q = S(16), S(19)           # x, y, z each (G1 coordinates embedded as (v, 0))
PA, QA, QN, Q1, Q2 = S(22), S(24), S(26), S(28), S(30)     # affine points (x, y)
RX, RY, RZ = S(32), S(33), S(34)    # the running point of the line walk
L0, LVW, LVV = S(35), S(36), S(37)  # one line: ell_0, ell_vw, ell_vv (field order of src/groups/mod.rs:566-571)
X2, X4 = S(38), S(39)               # ell_vv * P.x, ell_vw * P.y
F = S(40)                           # Fq12 accumulator / result, six slots c0.c0, c0.c1, c0.c2, c1.c0, c1.c1, c1.c2
TMP = S(48)                         # 48 temporaries
E_BASE, N_EREG = 96, 13             # Fq12 registers of the final exponentiation, six slots each
N_SLOTS = E_BASE + 6 * N_EREG       # 174


def t_(i):
    assert 0 <= i < 48
    return TMP + i


def ereg(k):
    assert 0 <= k < N_EREG
    return E_BASE + 6 * k


def fq6_products(dst, xs, ys):
    """Karatsuba lanes of an Fq6 product (src/fields/fq6.rs:194-208): dst + 0..5 = v0, v1, v2, m01, m12, m02."""
    xs, ys = [terms(x) for x in xs], [terms(y) for y in ys]
    return [(dst + 0, xs[0], ys[0]), (dst + 1, xs[1], ys[1]), (dst + 2, xs[2], ys[2]),
            (dst + 3, xs[0] + xs[1], ys[0] + ys[1]), (dst + 4, xs[1] + xs[2], ys[1] + ys[2]), (dst + 5, xs[0] + xs[2], ys[0] + ys[2])]


def fq6_outputs(p):
    """The three coefficients of the product from the six Karatsuba products at p + 0..5 (as term lists)."""
    v0, v1, v2, m01, m12, m02 = (T(p + i) for i in range(6))
    return [[v0, xi(m12), xi(neg(v1)), xi(neg(v2))],
            [m01, neg(v0), neg(v1), xi(v2)],
            [m02, neg(v0), neg(v2), v1]]


# ------------------------------------------------------------------------------------------------------------------
# routines
# ------------------------------------------------------------------------------------------------------------------
def build_routines():
    R = {}

    def new(name):
        r = Routine(name)
        R[name] = r
        return r

    # ---- joint affine conversion with ONE inversion (pair_to_affine of csrc/bn254_kernels.cuh; to_affine src/groups/mod.rs:198-217)
    r = new("TOAFF")
    px, py, pz = IN_P, IN_P + 1, IN_P + 2
    qx, qy, qz = IN_Q, IN_Q + 1, IN_Q + 2
    r.mul([(t_(0), qz, qz, False, True)])                        # norm(q.z)
    r.mul([(t_(1), pz, t_(0))])
    r.inv(t_(2), t_(1))
    r.mul([(t_(3), t_(2), t_(0)), (t_(4), t_(2), pz)])           # 1/p.z, 1/norm
    r.mul([(t_(5), t_(3), t_(3)), (t_(6), qz, t_(4), True, False)])   # 1/p.z^2, 1/q.z = conj(q.z)/norm
    r.mul([(PA, px, t_(5)), (t_(7), t_(5), t_(3)), (t_(8), t_(6), t_(6))])
    r.mul([(PA + 1, py, t_(7)), (QA, qx, t_(8)), (t_(9), t_(8), t_(6))])
    r.mul([(QA + 1, qy, t_(9))])
    # line walk starts at R = (Q.x, Q.y, 1); -Q for the NAF digits 3; f = 1
    r.lin([(RX, QA), (RY, QA + 1), (RZ, ONE), (QN, QA), (QN + 1, T(QA + 1, -1)), (F, ONE)] + [(F + i, ZERO) for i in range(1, 6)])

    # ---- doubling step (src/groups/mod.rs:754-776): R in place, line into L
    r = new("DBL")
    r.mul([(t_(0), RX, RY), (LVV, [RX, RX, RX], RX), (t_(2), RY, RY), (t_(3), RZ, RZ), (t_(4), [RY, RZ], [RY, RZ])])   # ell_vv = 3 X^2
    r.lin([(t_(5), t_(0), True),                                                   # a = X Y / 2
           (t_(6), [T(t_(4)), T(t_(2), -1), T(t_(3), -1)]),                        # h = (Y + Z)^2 - b - c
           (LVW, [T(t_(4), -1), T(t_(2)), T(t_(3))])])                             # ell_vw = -h
    r.mul([(t_(7), C_B3, t_(3))])                                                  # e = 3 b' Z^2
    r.lin([(L0, [T(t_(7), 1, 1), T(t_(2), -1, 1)]),                                # ell_0 = xi (e - b)
           (t_(8), [T(t_(2))] + times(T(t_(7)), 3), True),                         # g = (b + 3 e) / 2
           (t_(9), [T(t_(2))] + times(T(t_(7), -1), 3))])                          # b - f
    r.mul([(RX, t_(5), t_(9)), (t_(10), t_(8), t_(8)), (t_(11), [t_(7), t_(7), t_(7)], t_(7)), (RZ, t_(2), t_(6))])
    r.lin([(RY, [T(t_(10)), T(t_(11), -1)])])                                      # Y' = g^2 - 3 e^2

    # ---- mixed addition step (src/groups/mod.rs:731-752): R += the affine point at base B (x, y); line into L
    r = new("ADD")
    bx, by = B, B + 1
    r.mul([(t_(0), RZ, bx), (t_(1), RZ, by)])
    r.lin([(t_(2), [T(RX), T(t_(0), -1)]), (t_(3), [T(RY), T(t_(1), -1)]),          # d, e
           (LVW, [T(RX), T(t_(0), -1)]), (LVV, [T(RY, -1), T(t_(1))])])             # ell_vw = d, ell_vv = -e
    r.mul([(t_(4), t_(2), t_(2)), (t_(5), t_(3), t_(3)), (t_(6), t_(3), bx), (t_(7), t_(2), by)])   # f, g, e x, d y
    r.mul([(t_(8), t_(2), t_(4)), (t_(9), RX, t_(4)), (t_(10), RZ, t_(5))])         # h, i, Z g
    r.lin([(t_(11), [T(t_(10)), T(t_(8))] + times(T(t_(9), -1), 2)),                # j = Z g + h - 2 i
           (t_(12), times(T(t_(9)), 3) + [T(t_(10), -1), T(t_(8), -1)]),            # i - j
           (L0, [T(t_(6), 1, 1), T(t_(7), -1, 1)])])                                # ell_0 = xi (e x - d y)
    r.mul([(RX, t_(2), t_(11)), (t_(13), t_(3), t_(12)), (t_(14), t_(8), RY), (RZ, RZ, t_(8))])
    r.lin([(RY, [T(t_(13)), T(t_(14), -1)])])

    # ---- pi(Q), -pi^2(Q) (AffineG2::mul_by_q, src/groups/mod.rs:694-699, :718-726)
    r = new("FROBQ")
    r.mul([(Q1, QA, C_TWQX, True, False), (Q1 + 1, QA + 1, C_TWQY, True, False)])
    r.mul([(Q2, Q1, C_TWQX, True, False), (Q2 + 1, T(Q1 + 1, -1), C_TWQY, True, False)])

    # ---- f <- f * line(P) (src/groups/mod.rs:589 + Fq12::mul_by_024, src/fields/fq12.rs:130-196) as six 3-term sums
    r = new("LINE")
    r.mul([(X2, LVV, PA), (X4, LVW, PA + 1)])
    z = [F + i for i in range(6)]
    x0, x2, x4 = L0, X2, X4
    prods = [(z[0], x0), (z[1], x2), (z[4], x4),      # z0' = z0 x0 + xi (z1 x2 + z4 x4)
             (z[1], x0), (z[5], x4), (z[2], x2),      # z1' = z1 x0 + xi (z5 x4 + z2 x2)
             (z[2], x0), (z[0], x2), (z[3], x4),      # z2' = z2 x0 + z0 x2 + z3 x4
             (z[3], x0), (z[2], x4), (z[4], x2),      # z3' = z3 x0 + xi (z2 x4 + z4 x2)
             (z[4], x0), (z[0], x4), (z[5], x2),      # z4' = z4 x0 + z0 x4 + xi z5 x2
             (z[5], x0), (z[3], x2), (z[1], x4)]      # z5' = z5 x0 + z3 x2 + z1 x4
    r.mul([(t_(i), a, b) for i, (a, b) in enumerate(prods)])
    p = [T(t_(i)) for i in range(18)]
    r.lin([(z[0], [p[0], xi(p[1]), xi(p[2])]), (z[1], [p[3], xi(p[4]), xi(p[5])]), (z[2], [p[6], p[7], p[8]]),
           (z[3], [p[9], xi(p[10]), xi(p[11])]), (z[4], [p[12], p[13], xi(p[14])]), (z[5], [p[15], p[16], p[17]])])

    # ---- f <- f^2, complex method (src/fields/fq12.rs:295-303)
    r = new("SQR12")
    r.lin([(t_(0), [T(z[0]), xi(T(z[5]))]), (t_(1), [z[1], z[3]]), (t_(2), [z[2], z[4]])])      # a0 + v a1
    r.mul(fq6_products(t_(3), [z[0], z[1], z[2]], [z[3], z[4], z[5]]) +                          # ab = a0 a1
          fq6_products(t_(9), [[z[0], z[3]], [z[1], z[4]], [z[2], z[5]]], [t_(0), t_(1), t_(2)]))   # (a0 + a1)(a0 + v a1)
    oa, ot = fq6_outputs(t_(3)), fq6_outputs(t_(9))
    r.lin([(t_(15 + i), oa[i]) for i in range(3)] + [(t_(18 + i), ot[i]) for i in range(3)])
    ab, tt = [T(t_(15 + i)) for i in range(3)], [T(t_(18 + i)) for i in range(3)]
    r.lin([(z[0], [tt[0], neg(ab[0]), xi(neg(ab[2]))]), (z[1], [tt[1], neg(ab[1]), neg(ab[0])]), (z[2], [tt[2], neg(ab[2]), neg(ab[1])]),
           (z[3], times(ab[0], 2)), (z[4], times(ab[1], 2)), (z[5], times(ab[2], 2))])

    # ---- D <- A * B and D <- A * conj(B) (src/fields/fq12.rs:316-328; conj = unitary_inverse :126-128).  D may alias A or B.
    for name, s in (("MUL12", 1), ("MUL12C", -1)):
        r = new(name)
        a, b = [A + i for i in range(6)], [B + i for i in range(6)]
        r.mul(fq6_products(t_(0), a[0:3], b[0:3]) +
              fq6_products(t_(6), a[3:6], [T(b[3 + i], s) for i in range(3)]) +
              fq6_products(t_(12), [[a[i], a[3 + i]] for i in range(3)], [[T(b[i]), T(b[3 + i], s)] for i in range(3)]))
        outs = fq6_outputs(t_(0)) + fq6_outputs(t_(6)) + fq6_outputs(t_(12))
        r.lin([(t_(18 + k), outs[k]) for k in range(9)])
        aa, bb, cc = ([T(t_(18 + 3 * k + i)) for i in range(3)] for k in range(3))
        r.lin([(D + 0, [aa[0], xi(bb[2])]), (D + 1, [aa[1], bb[0]]), (D + 2, [aa[2], bb[1]])] +
              [(D + 3 + i, [cc[i], neg(aa[i]), neg(bb[i])]) for i in range(3)])

    # ---- D <- cyclotomic_squared(A) (src/fields/fq12.rs:198-247): three Fq4 squarings as a^2, b^2, (a + b)^2.  In place allowed.
    r = new("CYC")
    a = [A + i for i in range(6)]
    pairs = [(a[0], a[4]), (a[3], a[2]), (a[1], a[5])]          # (z0, z1), (z2, z3), (z4, z5) in the reference's labels
    lanes = []
    for k, (u, v) in enumerate(pairs):           # 3 a^2, 3 b^2, 6 a b: the factors 3 and 2 ride on the operand sums
        lanes += [(t_(3 * k), [u, u, u], u), (t_(3 * k + 1), [v, v, v], v), (t_(3 * k + 2), [u, u, u], [v, v])]
    r.mul(lanes)

    def t0(k, old):      # 3 (a^2 + xi b^2) - 2 old
        return [T(t_(3 * k)), T(old, -1), T(old, -1), T(t_(3 * k + 1), 1, 1)]

    def t1(k, old, with_xi):   # 3 (2ab) + 2 old, times xi for the last pair
        return [T(t_(3 * k + 2), 1, 1 if with_xi else 0), T(old), T(old)]
    r.lin([(D + 0, t0(0, a[0])), (D + 4, t1(0, a[4], False)),
           (D + 1, t0(1, a[1])), (D + 5, t1(1, a[5], False)),
           (D + 2, t0(2, a[2])), (D + 3, t1(2, a[3], True))])

    # ---- D <- conj(A), D <- A
    r = new("CONJ12")
    r.lin([(D + i, T(A + i, 1 if i < 3 else -1)) for i in range(6)])
    r = new("COPY12")
    r.lin([(D + i, A + i) for i in range(6)])

    # ---- D <- frobenius_map(A, p) (src/fields/fq12.rs:112-119, fq6.rs:125-131).  In place allowed.
    for p in (1, 2, 3):
        r = new("FROB%d" % p)
        cj = bool(p & 1)
        r.mul([(D + 0, A + 0, ONE, cj, False), (D + 1, A + 1, C_F6C1[p], cj, False), (D + 2, A + 2, C_F6C2[p], cj, False),
               (t_(0), A + 3, ONE, cj, False), (t_(1), A + 4, C_F6C1[p], cj, False), (t_(2), A + 5, C_F6C2[p], cj, False)])
        r.mul([(D + 3 + i, t_(i), C_F12C[p]) for i in range(3)])

    # ---- D <- A^-1 (src/fields/fq12.rs:305-313 -> fq6.rs:179-191 -> fq2.rs:119-130; the Fq inversion is Fermat, same element)
    r = new("INV12")
    a = [A + i for i in range(6)]
    r.mul(fq6_products(t_(0), a[0:3], a[0:3]) + fq6_products(t_(6), a[3:6], a[3:6]))
    o0, o1 = fq6_outputs(t_(0)), fq6_outputs(t_(6))
    r.lin([(t_(12 + i), o0[i]) for i in range(3)] + [(t_(15 + i), o1[i]) for i in range(3)])
    r.lin([(t_(18), [T(t_(12)), xi(T(t_(17), -1))]), (t_(19), [T(t_(13)), T(t_(15), -1)]), (t_(20), [T(t_(14)), T(t_(16), -1)])])
    g0, g1, g2 = t_(18), t_(19), t_(20)                           # t = a0^2 - v a1^2
    r.mul([(t_(21), g0, g0), (t_(22), g1, g2), (t_(23), g2, g2), (t_(24), g0, g1), (t_(25), g1, g1), (t_(26), g0, g2)])
    r.lin([(t_(27), [T(t_(21)), xi(T(t_(22), -1))]), (t_(28), [xi(T(t_(23))), T(t_(24), -1)]), (t_(29), [T(t_(25)), T(t_(26), -1)])])
    r.mul([(t_(30), g2, t_(28)), (t_(31), g1, t_(29)), (t_(32), g0, t_(27))])
    r.lin([(t_(33), [T(t_(32)), xi(T(t_(30))), xi(T(t_(31)))])])
    r.mul([(t_(34), t_(33), t_(33), False, True)])                # norm
    r.inv(t_(35), t_(34))
    r.mul([(t_(36), t_(33), t_(35), True, False)])                # conj / norm
    r.mul([(t_(37 + i), t_(36), t_(27 + i)) for i in range(3)])   # u = t^-1 (Fq6)
    u = [t_(37), t_(38), t_(39)]
    r.mul(fq6_products(t_(0), a[0:3], u) + fq6_products(t_(6), a[3:6], u))
    o0, o1 = fq6_outputs(t_(0)), fq6_outputs(t_(6))
    r.lin([(D + i, o0[i]) for i in range(3)] + [(D + 3 + i, [neg(t) for t in o1[i]]) for i in range(3)])
    build_paired_routines(new, R)
    return R


# ------------------------------------------------------------------------------------------------------------------
# paired routines of the quad-lane executor (csrc/bn_coop4.cuh): the G2 line walk on lanes 24..31 (one warp of the group) WHILE
# lanes 0..23 square f and multiply the previous lines into it.  The walk never reads f, so it runs one stage ahead; lines cross
# over through four buffers (two per stage, stages alternate).  During the Miller loop the Fq12 registers of the final
# exponentiation are idle: the walk's temporaries and the line buffers live there.  Appended AFTER the routines above, so the
# step numbers of the one-warp executor do not move.
# ------------------------------------------------------------------------------------------------------------------
def w_(i):
    """temporaries of the walk"""
    assert 0 <= i < 24
    return S(E_BASE + i)


YP, YN = S(E_BASE + 24), S(E_BASE + 25)      # the running point's Y as YP - YN (saves the walk a LIN step per doubling)
T0 = TMP + 47                                # z0 + xi z5 of the CURRENT Fq12 accumulator (first operand sum of the next squaring)
LINE_BUF = [E_BASE + 30 + 3 * j for j in range(4)]     # (ell_0, ell_vv P.x, ell_vw P.y) x [stage parity][doubling, addition]
PAIRED_W_LANES, PAIRED_F_LANES = 8, 24


def build_paired_routines(new, R):
    ry = [T(YP), T(YN, -1)]
    # ---- WINIT: the walk starts at R = (Q.x, Q.y, 1) with Y = YP - YN
    r = new("WINIT")
    r.lin([(YP, QA + 1), (YN, ZERO)])

    # ---- WDBL: doubling step (src/groups/mod.rs:754-776) + the two scalings of src/groups/mod.rs:589; the line goes to base D
    r = new("WDBL")
    r.mul([(w_(0), RX, ry), (w_(1), [RX, RX, RX], RX), (w_(2), ry, ry), (w_(3), RZ, RZ), (w_(4), ry + [T(RZ)], ry + [T(RZ)])])
    h = [T(w_(4)), T(w_(2), -1), T(w_(3), -1)]                                   # h = (Y + Z)^2 - b - c
    r.mul([(w_(7), C_B3, w_(3)),                                                 # e = 3 b' Z^2
           (D + 1, w_(1), PA),                                                   # ell_vv P.x, ell_vv = 3 X^2
           (D + 2, [neg(t) for t in h], PA + 1),                                 # ell_vw P.y, ell_vw = -h
           (RZ, w_(2), h)])                                                      # Z' = b h
    r.lin([(D + 0, [T(w_(7), 1, 1), T(w_(2), -1, 1)]),                           # ell_0 = xi (e - b)
           (w_(5), w_(0), True),                                                 # a = X Y / 2
           (w_(8), [T(w_(2))] + times(T(w_(7)), 3), True)])                      # g = (b + 3 e) / 2
    r.mul([(RX, w_(5), [T(w_(2))] + times(T(w_(7), -1), 3)),                     # X' = a (b - 3 e)
           (YP, w_(8), w_(8)), (YN, [w_(7), w_(7), w_(7)], w_(7))])              # Y' = g^2 - 3 e^2

    # ---- WADD: mixed addition step (src/groups/mod.rs:731-752) with the affine point at base B, line to base D
    r = new("WADD")
    bx, by = B, B + 1
    r.mul([(w_(10), RZ, bx), (w_(11), RZ, by)])
    d = [T(RX), T(w_(10), -1)]                                                   # d = X - Z x
    e = ry + [T(w_(11), -1)]                                                     # e = Y - Z y
    r.mul([(w_(12), d, d), (w_(13), e, e), (w_(14), e, bx), (w_(15), d, by),     # f, g, e x, d y
           (D + 1, [neg(t) for t in e], PA), (D + 2, d, PA + 1)])                # ell_vv = -e, ell_vw = d, scaled
    r.mul([(w_(16), d, w_(12)), (w_(17), RX, w_(12)), (w_(18), RZ, w_(13))])     # h, i, Z g
    r.lin([(w_(19), [T(w_(18)), T(w_(16))] + times(T(w_(17), -1), 2)),           # j = Z g + h - 2 i
           (w_(20), times(T(w_(17)), 3) + [T(w_(18), -1), T(w_(16), -1)]),       # i - j
           (D + 0, [T(w_(14), 1, 1), T(w_(15), -1, 1)])])                        # ell_0 = xi (e x - d y)
    r.mul([(w_(21), d, w_(19)), (w_(22), e, w_(20)), (w_(23), w_(16), ry), (RZ, RZ, w_(16))])
    r.lin([(RX, w_(21)), (YP, [T(w_(22)), T(w_(23), -1)]), (YN, ZERO)])

    # ---- LINEB: f <- f * line with (x0, x2, x4) already scaled at base A; also T0 = z0 + xi z5 of the new f
    r = new("LINEB")
    z = [F + i for i in range(6)]
    x0, x2, x4 = A + 0, A + 1, A + 2
    prods = [(z[0], x0), (z[1], x2), (z[4], x4), (z[1], x0), (z[5], x4), (z[2], x2), (z[2], x0), (z[0], x2), (z[3], x4),
             (z[3], x0), (z[2], x4), (z[4], x2), (z[4], x0), (z[0], x4), (z[5], x2), (z[5], x0), (z[3], x2), (z[1], x4)]
    r.mul([(t_(i), a, b) for i, (a, b) in enumerate(prods)])
    p = [T(t_(i)) for i in range(18)]
    r.lin([(z[0], [p[0], xi(p[1]), xi(p[2])]), (z[1], [p[3], xi(p[4]), xi(p[5])]), (z[2], [p[6], p[7], p[8]]),
           (z[3], [p[9], xi(p[10]), xi(p[11])]), (z[4], [p[12], p[13], xi(p[14])]), (z[5], [p[15], p[16], p[17]]),
           (T0, [p[0], xi(p[1]), xi(p[2]), xi(p[15]), xi(p[16]), xi(p[17])])])

    # ---- SQR12B: f <- f^2 (src/fields/fq12.rs:295-303) with a0 + v a1 taken as operand sums (T0 from LINEB): one LIN step less
    r = new("SQR12B")
    r.mul(fq6_products(t_(3), [z[0], z[1], z[2]], [z[3], z[4], z[5]]) +
          fq6_products(t_(9), [[z[0], z[3]], [z[1], z[4]], [z[2], z[5]]], [T0, [z[1], z[3]], [z[2], z[4]]]))
    oa, ot = fq6_outputs(t_(3)), fq6_outputs(t_(9))
    r.lin([(t_(15 + i), oa[i]) for i in range(3)] + [(t_(18 + i), ot[i]) for i in range(3)])
    ab, tt = [T(t_(15 + i)) for i in range(3)], [T(t_(18 + i)) for i in range(3)]
    r.lin([(z[0], [tt[0], neg(ab[0]), xi(neg(ab[2]))]), (z[1], [tt[1], neg(ab[1]), neg(ab[0])]), (z[2], [tt[2], neg(ab[2]), neg(ab[1])]),
           (z[3], times(ab[0], 2)), (z[4], times(ab[1], 2)), (z[5], times(ab[2], 2)),
           (T0, [tt[0], neg(ab[0]), xi(ab[2])])])                      # z0' + xi z5' (a squaring may follow directly: Gt::pow)
    # ---- Gt::pow of the quad-lane executor (k_coop4_gt_pow): SQR12B on the accumulator in F, then MUL12T = MUL12 whose last
    # LIN step also leaves T0 = d0 + xi d5 for the squaring that follows; T0INIT sets it for a value copied into F.
    r = new("MUL12T")
    a, b = [A + i for i in range(6)], [B + i for i in range(6)]
    r.mul(fq6_products(t_(0), a[0:3], b[0:3]) + fq6_products(t_(6), a[3:6], b[3:6]) +
          fq6_products(t_(12), [[a[i], a[3 + i]] for i in range(3)], [[b[i], b[3 + i]] for i in range(3)]))
    outs = fq6_outputs(t_(0)) + fq6_outputs(t_(6)) + fq6_outputs(t_(12))
    r.lin([(t_(18 + k), outs[k]) for k in range(9)])
    aa, bb, cc = ([T(t_(18 + 3 * k + i)) for i in range(3)] for k in range(3))
    r.lin([(D + 0, [aa[0], xi(bb[2])]), (D + 1, [aa[1], bb[0]]), (D + 2, [aa[2], bb[1]])] +
          [(D + 3 + i, [cc[i], neg(aa[i]), neg(bb[i])]) for i in range(3)] +
          [(T0, [aa[0], xi(cc[2]), xi(neg(aa[2]))])])                  # d0 + xi d5 = aa0 + xi bb2 + xi (cc2 - aa2 - bb2)
    r = new("T0INIT")
    r.lin([(T0, [T(F), T(F + 5, 1, 1)])])
    for name in ("WINIT", "WDBL", "WADD"):
        assert all(len(st.lanes) <= PAIRED_W_LANES for st in R[name].steps), name
    for name in ("LINEB", "SQR12B"):
        assert all(len(st.lanes) <= PAIRED_F_LANES for st in R[name].steps), name


# ------------------------------------------------------------------------------------------------------------------
# programs (call lists)
# ------------------------------------------------------------------------------------------------------------------
def wnaf4(k):
    """Width-4 signed-window digits of k, least significant first: odd digits in -7..7, at most one in any 4 positions."""
    d = []
    while k:
        if k & 1:
            z = k & 15
            if z > 8:
                z -= 16
            k -= z
        else:
            z = 0
        d.append(z)
        k >>= 1
    return d


def build_program():
    """Returns (calls, marks): calls = [(routine, baseD, baseA, baseB)], marks = index ranges of the phases."""
    calls = []

    def call(name, d=0, a=0, b=0):
        calls.append((name, d, a, b))

    marks = {}
    # ---- Miller loop with the line coefficients produced on the fly (AffineG2::precompute + G2Precomp::miller_loop)
    marks["pairing_begin"] = len(calls)
    call("TOAFF")
    marks["miller_begin"] = len(calls)
    naf = o.ATE_LOOP_COUNT_NAF
    assert len(naf) == 64
    for k, dg in enumerate(naf):
        if k:
            call("SQR12")                   # the reference squares 1 in the first iteration: same value
        call("DBL")
        call("LINE")
        if dg:
            call("ADD", 0, 0, (QA if dg == 1 else QN).off)
            call("LINE")
    call("FROBQ")
    call("ADD", 0, 0, Q1.off)
    call("LINE")
    call("ADD", 0, 0, Q2.off)
    call("LINE")
    marks["miller_end"] = len(calls)
    # ---- final exponentiation (src/fields/fq12.rs:62-110) on the Fq12 registers
    marks["fexp_begin"] = len(calls)
    X, RA, RB, RC, RD, RE, RF, RG, CUR, A3, A5, A7, TT = (ereg(k) for k in range(13))
    f = F.off
    # easy part: x = frob2(c) c, c = conj(f) f^-1
    call("INV12", RB, f)
    call("MUL12C", RC, RB, f)               # c = f^-1 conj(f)
    call("FROB2", RA, RC)
    call("MUL12", X, RA, RC)
    marks["hard_begin"] = len(calls)

    digits = wnaf4(o.BN_U)
    assert sum(dg << i for i, dg in enumerate(digits)) == o.BN_U and digits[-1] == 1
    tab = {3: A3, 5: A5, 7: A7}

    def exp_by_neg_z(dst, a):
        """dst = conj(a^u), width-4 signed window (csrc/bn_tower.cuh fq12_exp_by_neg_z; value of src/fields/fq12.rs:121-124)."""
        call("CYC", CUR, a)                 # a^2: serves the table and, the top digit being +1, the first doubling
        call("MUL12", A3, CUR, a)
        call("MUL12", A5, CUR, A3)
        call("MUL12", A7, CUR, A5)
        tab[1] = a
        first = True
        for i in range(len(digits) - 2, -1, -1):
            if not first:
                call("CYC", CUR, CUR)
            first = False
            dg = digits[i]
            if dg:
                call("MUL12" if dg > 0 else "MUL12C", CUR, CUR, tab[abs(dg)])
        call("CONJ12", dst, CUR)

    exp_by_neg_z(RA, X)                     # a
    call("CYC", RB, RA)                     # b
    call("CYC", RC, RB)                     # c
    call("MUL12", RD, RC, RB)               # d
    exp_by_neg_z(RE, RD)                    # e
    call("CYC", RF, RE)                     # f
    exp_by_neg_z(RG, RF)                    # g
    call("MUL12C", RA, RE, RG)              # j = conj(g) e
    call("MUL12C", RC, RA, RD)              # k = j conj(d)
    call("MUL12", RF, RC, RB)               # l = k b
    call("MUL12", RA, RC, RE)               # m = k e
    call("MUL12", RG, X, RA)                # n = x m
    call("FROB1", RA, RF)                   # o
    call("MUL12", RB, RA, RG)               # p = o n
    call("FROB2", RA, RC)                   # q
    call("MUL12", RG, RA, RB)               # r = q p
    call("MUL12C", RA, RF, X)               # t = conj(x) l
    call("FROB3", RB, RA)                   # u
    call("MUL12", f, RB, RG)                # v = u r  -> the result slots
    marks["fexp_end"] = len(calls)
    marks["pairing_end"] = len(calls)
    return calls, marks


def build_paired_miller():
    """The Miller loop of the quad-lane executor as stages (f_calls, w_calls): f_calls = [(routine, base A)] run on lanes 0..23,
    w_calls = [(routine, base D, base B)] on lanes 24..31 at the same time; a stage ends when both lists are done.  Walk stage k
    (the line steps of NAF digit k; stages 64, 65: the additions of pi(Q), -pi^2(Q)) runs beside f stage k - 1."""
    naf = o.ATE_LOOP_COUNT_NAF
    walk = []                                   # per walk stage: [(routine, D, B)] and the buffers it fills
    for k, dg in enumerate(naf):
        par = k & 1
        calls = [("WDBL", LINE_BUF[2 * par], 0)]
        if dg:
            calls.append(("WADD", LINE_BUF[2 * par + 1], (QA if dg == 1 else QN).off))
        walk.append(calls)
    walk.append([("WADD", LINE_BUF[0], Q1.off)])                 # stage 64
    walk.append([("WADD", LINE_BUF[2], Q2.off)])                 # stage 65
    fch = []
    for k in range(len(walk)):
        calls = [("SQR12B", 0)] if 0 < k < 64 else []            # the reference squares 1 in the first iteration: same value
        calls += [("LINEB", d) for _, d, _ in walk[k]]
        fch.append(calls)
    stages = [([], [("WINIT", 0, 0)] + walk[0])]
    for k in range(len(walk)):
        stages.append((fch[k], walk[k + 1] if k + 1 < len(walk) else []))
    return stages


def flatten_paired(R, index):
    """Trace of the quad-lane executor: entries (step F or None, step W or None, D, A, B).  TOAFF and FROBQ first (all lanes),
    then the paired Miller stages, then the final exponentiation (all lanes, the sequential program's calls)."""
    calls, marks = build_program()
    trace, tm = [], {}
    tm["pairing_begin"] = 0
    for name in ("TOAFF", "FROBQ"):
        for k in range(len(R[name].steps)):
            trace.append((index[(name, k)], None, 0, 0, 0))
    for f_calls, w_calls in build_paired_miller():
        fs = [(index[(n, k)], a) for n, a in f_calls for k in range(len(R[n].steps))]
        ws = [(index[(n, k)], d, b) for n, d, b in w_calls for k in range(len(R[n].steps))]
        for i in range(max(len(fs), len(ws))):
            sf, a = fs[i] if i < len(fs) else (None, 0)
            sw, d, b = ws[i] if i < len(ws) else (None, 0, 0)
            trace.append((sf, sw, d, a, b))
    tm["miller_end"] = tm["fexp_begin"] = len(trace)
    for name, bd, ba, bb in calls[marks["fexp_begin"]:marks["fexp_end"]]:
        for k in range(len(R[name].steps)):
            trace.append((index[(name, k)], None, bd, ba, bb))
    tm["fexp_end"] = tm["pairing_end"] = len(trace)
    return trace, tm


def run_paired_entry(slots, steps, entry):
    """One trace entry of the quad-lane executor on the simulator: both steps read the slots as they were, then write."""
    sf, sw, d, a, b = entry
    bases = (d, a, b)
    if sw is None:
        run_step(slots, steps[sf], bases)
        return
    assert sf is None or len(steps[sf].lanes) <= PAIRED_F_LANES
    assert len(steps[sw].lanes) <= PAIRED_W_LANES
    before = list(slots)
    merged = {}
    touched = []
    for sid in (sf, sw):
        if sid is None:
            continue
        work = list(before)
        run_step(work, steps[sid], bases)
        reads = set()
        for ln in steps[sid].lanes:
            for k in ("a", "b", "t"):
                for t in ln.get(k, []):
                    reads.add(_resolve(t[0], bases))
        writes = {i for i in range(len(work)) if work[i] is not before[i]}
        writes |= {_resolve(ln["dst"], bases) for ln in steps[sid].lanes}
        touched.append((reads, writes))
        for i in writes:
            assert i not in merged, "both halves of a paired step write slot %d" % i
            merged[i] = work[i]
    if len(touched) == 2:
        (r0, w0), (r1, w1) = touched
        assert not (w0 & r1) and not (w1 & r0), "one half of a paired step reads what the other writes: %s %s" % (w0 & r1, w1 & r0)
    for i, v in merged.items():
        slots[i] = v


def simulate_paired(p, q, phase="pairing"):
    """The quad-lane executor's program on Jacobian points (canonical integer tuples), from the routines."""
    R = build_routines()
    index, steps = {}, []
    for name, r in R.items():
        for k, st in enumerate(r.steps):
            index[(name, k)] = len(steps)
            steps.append(st)
    trace, tm = flatten_paired(R, index)
    slots = fresh_slots()
    for i in range(3):
        slots[IN_P.off + i] = (p[i] % Q, 0)
        slots[IN_Q.off + i] = q[i]
    for entry in trace[:tm["miller_end"] if phase == "miller" else tm["pairing_end"]]:
        run_paired_entry(slots, steps, entry)
    return get_f12(slots, F.off)


def constants():
    """Values of the constant slots (canonical Fq2 tuples)."""
    c = {ZERO.off: (0, 0), ONE.off: (1, 0), C_B3.off: o.fq2_scale(o.G2_B, 3), C_TWQX.off: o.TWIST_MUL_BY_Q_X, C_TWQY.off: o.TWIST_MUL_BY_Q_Y}
    for p in (1, 2, 3):
        c[C_F6C1[p].off] = o.FQ6_FROB_C1[p]
        c[C_F6C2[p].off] = o.FQ6_FROB_C2[p]
        c[C_F12C[p].off] = o.FQ12_FROB_C1[p]
    return c


def fresh_slots():
    slots = [(0, 0)] * N_SLOTS
    for k, v in constants().items():
        slots[k] = v
    return slots


def get_f12(slots, base):
    s = slots[base:base + 6]
    return ((s[0], s[1], s[2]), (s[3], s[4], s[5]))


def put_f12(slots, base, v):
    for i in range(6):
        slots[base + i] = v[i // 3][i % 3]


def simulate(p, q, phase="pairing"):
    """Run the program on Jacobian points p (G1) and q (G2) given as canonical integer tuples.  Returns the Fq12 in the result
    slots (phase 'miller': the un-exponentiated value)."""
    R = build_routines()
    calls, marks = build_program()
    slots = fresh_slots()
    for i in range(3):
        slots[IN_P.off + i] = (p[i] % Q, 0)
        slots[IN_Q.off + i] = q[i]
    lo = marks["pairing_begin"]
    hi = marks["miller_end"] if phase == "miller" else marks["pairing_end"]
    run_program(slots, R, calls[lo:hi])
    return get_f12(slots, F.off)


def simulate_pow(a, k):
    """Gt::pow the way k_coop_gt_pow walks it: MSB first, SQR12 on the accumulator in F, MUL12 (D = A = F, B = the base in
    register 1) on set bits; leading zero bits skipped (FieldElement::pow squares one there: same value)."""
    R = build_routines()
    slots = fresh_slots()
    put_f12(slots, ereg(1), a)
    put_f12(slots, F.off, o.FQ12_ONE)
    started = False
    for i in range(255, -1, -1):
        if started:
            run_program(slots, R, [("SQR12", 0, 0, 0)])
        if (k >> i) & 1:
            if started:
                run_program(slots, R, [("MUL12", F.off, F.off, ereg(1))])
            else:
                put_f12(slots, F.off, a)
                started = True
    return get_f12(slots, F.off)


def simulate_pow4(a, k):
    """Gt::pow the way k_coop4_gt_pow walks it: T0INIT after the copy of the base, then SQR12B / MUL12T."""
    R = build_routines()
    slots = fresh_slots()
    put_f12(slots, ereg(1), a)
    put_f12(slots, F.off, o.FQ12_ONE)
    started = False
    for i in range(255, -1, -1):
        if started:
            run_program(slots, R, [("SQR12B", 0, 0, 0)])
        if (k >> i) & 1:
            if started:
                run_program(slots, R, [("MUL12T", F.off, F.off, ereg(1))])
            else:
                put_f12(slots, F.off, a)
                run_program(slots, R, [("T0INIT", 0, 0, 0)])
                started = True
    return get_f12(slots, F.off)


def simulate_fexp(x):
    R = build_routines()
    calls, marks = build_program()
    slots = fresh_slots()
    put_f12(slots, F.off, x)
    run_program(slots, R, calls[marks["fexp_begin"]:marks["fexp_end"]])
    return get_f12(slots, F.off)


# ------------------------------------------------------------------------------------------------------------------
# emitter: csrc/bn_coop_ucode.cuh
# ------------------------------------------------------------------------------------------------------------------
ZERO_TERM = 0        # slot 0 (the constant zero), plus sign


def encode_term(t):
    ref, sign, _ = t
    return ref.code() | ((1 if sign < 0 else 0) << 10)


def encode_step(st):
    """header word and 5 x 32 lane words.  Lane word 0: dst | active << 10 | conjA << 11 | conjB << 12 | half << 13; words 1..4: eight
    16-bit terms (slot code | minus << 10), first n0 terms = A side / group 0, next n1 = B side / group 1."""
    n0, n1 = st.counts()
    words = [[0] * 32 for _ in range(5)]
    for lane in range(32):
        if lane < len(st.lanes):
            ln = st.lanes[lane]
            if st.op == OP_MUL:
                s0, s1 = ln["a"], ln["b"]
            elif st.op == OP_LIN:
                s0, s1 = [t for t in ln["t"] if t[2] == 0], [t for t in ln["t"] if t[2] == 1]
            else:
                s0, s1 = ln["t"], []
            tl = [encode_term(t) for t in s0] + [ZERO_TERM] * (n0 - len(s0)) + [encode_term(t) for t in s1] + [ZERO_TERM] * (n1 - len(s1))
            tl += [ZERO_TERM] * (NT - len(tl))
            w0 = ln["dst"].code() | (1 << 10) | (int(ln.get("ca", False)) << 11) | (int(ln.get("cb", False)) << 12) | (int(ln.get("half", False)) << 13)
        else:
            tl, w0 = [ZERO_TERM] * NT, NO_DST
        words[0][lane] = w0
        for k in range(4):
            words[1 + k][lane] = tl[2 * k] | (tl[2 * k + 1] << 16)
    return st.header(), words


def flatten():
    """(steps, trace, marks): the distinct steps of all routines and the executed sequence (step index, bases) of the program."""
    R = build_routines()
    calls, marks = build_program()
    index, steps = {}, []
    for name, r in R.items():
        for k, st in enumerate(r.steps):
            index[(name, k)] = len(steps)
            steps.append(st)
    trace, tmarks, pos = [], {}, {}
    for ci, (name, bd, ba, bb) in enumerate(calls):
        pos[ci] = len(trace)
        for k in range(len(R[name].steps)):
            trace.append((index[(name, k)], bd, ba, bb))
    pos[len(calls)] = len(trace)
    for k, v in marks.items():
        tmarks[k] = pos[v]
    return steps, trace, tmarks


NONE_STEP = 0xff
N_SEQ_ROUTINE = "WINIT"          # the first routine the one-warp executor never runs: its tables end there


def flatten4():
    """(steps, paired trace, marks, number of steps the one-warp executor needs)."""
    R = build_routines()
    index, steps = {}, []
    for name, r in R.items():
        for k, st in enumerate(r.steps):
            index[(name, k)] = len(steps)
            steps.append(st)
    trace, tm = flatten_paired(R, index)
    return steps, trace, tm, index[(N_SEQ_ROUTINE, 0)]


def encode_paired(trace):
    """Two words per entry: step F | step W << 8 (0xff = none), then base D << 8 | base A << 16 | base B << 24."""
    words = []
    for sf, sw, d, a, b in trace:
        words.append((NONE_STEP if sf is None else sf) | ((NONE_STEP if sw is None else sw) << 8))
        words.append((d << 8) | (a << 16) | (b << 24))
    return words


def mont_limbs(v):
    m = v * o.MONT_R % Q
    return [(m >> (32 * i)) & 0xffffffff for i in range(8)]


def emit(path):
    steps, trace, marks = flatten()
    out = []
    w = out.append
    w("// GENERATED by tools/gen_coop_ucode.py -- do not edit.  Microcode of the warp-cooperative pairing (csrc/bn_coop.cuh):")
    w("// %d distinct steps, %d executed steps per pairing.  The simulator in the generator runs these very tables against the oracle." % (len(steps), len(trace)))
    w("#pragma once")
    w("#include <stdint.h>")
    w("#define BN_COOP_NSLOTS %d" % ((N_SLOTS + 1) & ~1))
    w("#define BN_COOP_NSTEPS %d" % len(steps))
    w("#define BN_COOP_NTRACE %d" % len(trace))
    w("#define BN_COOP_NCONST %d" % N_CONST)
    w("#define BN_COOP_SLOT_IN_P %d" % IN_P.off)
    w("#define BN_COOP_SLOT_IN_Q %d" % IN_Q.off)
    w("#define BN_COOP_SLOT_F %d" % F.off)
    for k in ("pairing_begin", "miller_end", "fexp_begin", "fexp_end", "pairing_end"):
        w("#define BN_COOP_T_%s %d" % (k.upper(), marks[k]))
    # step ranges of the routines a kernel may call on its own (Gt::pow walks its scalar with SQR12 / MUL12)
    R = build_routines()
    pos = 0
    for name, r in R.items():
        if name in ("SQR12", "MUL12", "SQR12B", "MUL12T", "T0INIT"):
            w("#define BN_COOP_R_%s_BEGIN %d" % (name, pos))
            w("#define BN_COOP_R_%s_COUNT %d" % (name, len(r.steps)))
        pos += len(r.steps)
    w("#define BN_COOP_SLOT_EREG1 %d" % ereg(1))
    hdrs, code = [], []
    for st in steps:
        h, words = encode_step(st)
        hdrs.append(h)
        for k in range(5):
            code.extend(words[k])

    def arr(name, vals, per=12):
        w("__device__ const uint32_t %s[%d] = {" % (name, len(vals)))
        for i in range(0, len(vals), per):
            w("    " + ", ".join("0x%08xu" % v for v in vals[i:i + per]) + ",")
        w("};")
    arr("bn_coop_hdr", hdrs)
    arr("bn_coop_code", code, 16)
    arr("bn_coop_trace", [s | (bd << 8) | (ba << 16) | (bb << 24) for s, bd, ba, bb in trace])
    # the quad-lane executor's trace (csrc/bn_coop4.cuh): paired steps, two words per entry
    steps4, trace4, marks4, nseq = flatten4()
    assert steps4 == steps or len(steps4) == len(steps)
    assert all(ts[0] < nseq for ts in trace), "the one-warp executor's trace stays inside its prefix of the step table"
    w("#define BN_COOP_NSTEPS_SEQ %d" % nseq)
    w("#define BN_COOP4_NTRACE %d" % len(trace4))
    for k in ("pairing_begin", "miller_end", "fexp_begin", "fexp_end", "pairing_end"):
        w("#define BN_COOP4_T_%s %d" % (k.upper(), marks4[k]))
    w("#define BN_COOP4_W_LANE0 %d" % PAIRED_F_LANES)
    arr("bn_coop4_trace", encode_paired(trace4))
    cs = constants()
    vals = []
    for k in range(N_CONST):
        c = cs.get(k, (0, 0))
        vals += mont_limbs(c[0]) + mont_limbs(c[1])
    arr("bn_coop_const", vals, 8)
    assert len(steps) < 256
    open(path, "w").write("\n".join(out) + "\n")
    return len(steps), len(trace)


def decode_and_run(p, q):
    """Simulate from the ENCODED tables (what the GPU executes): decodes bn_coop words back into terms."""
    steps, trace, marks = flatten()
    enc = [encode_step(st) for st in steps]
    slots = fresh_slots()
    for i in range(3):
        slots[IN_P.off + i] = (p[i] % Q, 0)
        slots[IN_Q.off + i] = q[i]
    half_inv = pow(2, -1, Q)
    for s, bd, ba, bb in trace[marks["pairing_begin"]:marks["pairing_end"]]:
        h, words = enc[s]
        op, n0, n1 = h & 15, (h >> 4) & 15, (h >> 8) & 15
        bases = (0, bd, ba, bb)
        writes = []
        for lane in range(32):
            w0 = words[0][lane]
            if not (w0 >> 10) & 1:
                continue
            tl = []
            for k in range(4):
                tl += [words[1 + k][lane] & 0xffff, words[1 + k][lane] >> 16]

            def acc(ts):
                a = (0, 0)
                for t in ts:
                    v = slots[(t & 0xff) + bases[(t >> 8) & 3]]
                    a = o.fq2_sub(a, v) if (t >> 10) & 1 else o.fq2_add(a, v)
                return a
            x, y = acc(tl[:n0]), acc(tl[n0:n0 + n1])
            if op == OP_MUL:
                if (w0 >> 11) & 1:
                    x = _conj(x)
                if (w0 >> 12) & 1:
                    y = _conj(y)
                r = o.fq2_mul(x, y)
            elif op == OP_LIN:
                r = o.fq2_add(x, o.fq2_mul_by_nonresidue(y))
                if (w0 >> 13) & 1:
                    r = (r[0] * half_inv % Q, r[1] * half_inv % Q)
            else:
                r = (pow(x[0], Q - 2, Q), 0)
            writes.append(((w0 & 0xff) + bases[(w0 >> 8) & 3], r))
        for d, r in writes:
            slots[d] = r
    return get_f12(slots, F.off)


def decode_and_run_paired(p, q):
    """The quad-lane executor's view of the ENCODED tables: per entry, lanes 24..31 take lanes 0..7 of step W when there is one,
    the other lanes (all 32 without a step W) take step F; reads see the slots as they were before the entry."""
    steps, trace, marks, _ = flatten4()
    enc = [encode_step(st) for st in steps]
    tw = encode_paired(trace)
    slots = fresh_slots()
    for i in range(3):
        slots[IN_P.off + i] = (p[i] % Q, 0)
        slots[IN_Q.off + i] = q[i]
    half_inv = pow(2, -1, Q)
    for e in range(marks["pairing_begin"], marks["pairing_end"]):
        w0, w1 = tw[2 * e], tw[2 * e + 1]
        sf, sw = w0 & 0xff, (w0 >> 8) & 0xff
        bases = (0, (w1 >> 8) & 0xff, (w1 >> 16) & 0xff, (w1 >> 24) & 0xff)
        writes = []
        for lane in range(32):
            if sw != NONE_STEP and lane >= PAIRED_F_LANES:
                sid, ln = sw, lane - PAIRED_F_LANES
            else:
                sid, ln = sf, lane
            if sid == NONE_STEP:
                continue
            h, words = enc[sid]
            op, n0, n1 = h & 15, (h >> 4) & 15, (h >> 8) & 15
            lw = words[0][ln]
            if not (lw >> 10) & 1:
                continue
            tl = []
            for k in range(4):
                tl += [words[1 + k][ln] & 0xffff, words[1 + k][ln] >> 16]

            def acc(ts):
                a = (0, 0)
                for t in ts:
                    v = slots[(t & 0xff) + bases[(t >> 8) & 3]]
                    a = o.fq2_sub(a, v) if (t >> 10) & 1 else o.fq2_add(a, v)
                return a
            x, y = acc(tl[:n0]), acc(tl[n0:n0 + n1])
            if op == OP_MUL:
                if (lw >> 11) & 1:
                    x = _conj(x)
                if (lw >> 12) & 1:
                    y = _conj(y)
                r = o.fq2_mul(x, y)
            elif op == OP_LIN:
                r = o.fq2_add(x, o.fq2_mul_by_nonresidue(y))
                if (lw >> 13) & 1:
                    r = (r[0] * half_inv % Q, r[1] * half_inv % Q)
            else:
                r = (pow(x[0], Q - 2, Q), 0)
            writes.append(((lw & 0xff) + bases[(lw >> 8) & 3], r))
        for d, r in writes:
            slots[d] = r
    return get_f12(slots, F.off)


if __name__ == "__main__":
    import random
    if "--check" in sys.argv:
        rng = random.Random(1)
        for _ in range(3):
            p = o.g_mul(o.FQ_OPS, o.G1_ONE, rng.randrange(1, o.R_ORDER))
            q = o.g_mul(o.FQ2_OPS, o.G2_ONE, rng.randrange(1, o.R_ORDER))
            assert simulate(p, q) == o.pairing(p, q) == decode_and_run(p, q) == simulate_paired(p, q) == decode_and_run_paired(p, q)
        print("microcode matches the oracle")
    else:
        n = emit(os.path.join(ROOT, "paritytech-bn_b200", "csrc", "bn_coop_ucode.cuh"))
        print("wrote bn_coop_ucode.cuh: %d steps, %d executed per pairing on one warp, %d on four" % (n + (len(flatten4()[1]),)))
