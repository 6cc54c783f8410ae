#!/usr/bin/env python3
"""Generate the PTX carry-chain blocks for 8x32-bit-limb BN254 base-field arithmetic.

Output: paritytech-bn_b200/csrc/fq_ptx.cuh (committed; regenerate with `python tools/gen_fq_ptx.py`).

Why a generator: there is no GPU in the build container, so the instruction sequences cannot be
executed here.  Each routine is therefore built once as an abstract instruction list (PTX
mnemonics over symbolic 32-bit registers + one carry flag), which is
  (1) SIMULATED in Python against big-int ground truth on random and edge inputs (--check), and
  (2) printed 1:1 as a single inline-asm block (one block per routine, so the compiler can never
      schedule a carry-clobbering instruction inside a chain).

Algorithm (our own derivation, see DESIGN.md "Fq multiply"): Montgomery multiplication a*b*2^-256
mod q, operand scanning, with the partial products split over two accumulators by the PARITY of
their limb position.  A 32x32->64 product at limb position p occupies limbs (p, p+1); products
whose positions share a parity never overlap, so the four products a0*b, a2*b, a4*b, a6*b form one
contiguous 8-limb chain `mad.lo.cc / madc.hi.cc / madc.lo.cc ...` with a single carry running
through it, and a1,a3,a5,a7 form a second chain one limb higher.  ptxas fuses each
(mad*.lo.cc, madc.hi.cc) pair into one IMAD.WIDE.U32[.X].  Reduction rows (k = t_i * (-q^-1) mod
2^32, t += k*q) are interleaved with the multiplication rows (CIOS) so the window stays 9 limbs.

Routines emitted:
  fq_mul_ptx      r = a*b/R mod q, fully reduced (in [0,q))
  fq_mulwide_ptx  t[16] = a*b                     (no reduction; for lazy reduction in Fq2)
  fq_redc_ptx     r = t/R mod q for t < q*R, fully reduced
  fq_add_ptx / fq_sub_ptx / fq_neg-free helpers: modular add/sub with the conditional correction
  wide_add_ptx / wide_sub_ptx: 512-bit add and subtract (sub adds q*2^256 back on borrow)
"""
import argparse
import os
import random
import sys

Q = 0x30644e72e131a029b85045b68181585d97816a916871ca8d3c208c16d87cfd47
M32 = (1 << 32) - 1
QL = [(Q >> (32 * i)) & M32 for i in range(8)]
INV32 = (-pow(Q, -1, 1 << 32)) % (1 << 32)
assert INV32 == 0xe4866389


class Prog:
    """Abstract PTX program: list of (op, dst, srcs).  Registers are strings; ints are immediates."""

    def __init__(self, name):
        self.name = name
        self.ins = []
        self.ntmp = 0
        self.inputs = []   # list of register names (order = asm operand order)
        self.outputs = []

    def tmp(self):
        self.ntmp += 1
        return "t%d" % (self.ntmp - 1)

    def emit(self, op, dst, *srcs, guard=None):
        """guard: None or a predicate name like "%p0" / "!%p1" (instruction executes only if true)."""
        self.ins.append((op, dst, srcs, guard))
        return dst

    def label(self, name):
        self.ins.append(("label", name, (), None))

    def bra(self, name, guard=None):
        self.ins.append(("bra", name, (), guard))

    # -- simulation -------------------------------------------------------------------------
    def run(self, env):
        env = dict(env)
        cf = 0
        pred = {}

        def val(s):
            return s if isinstance(s, int) else env[s]

        labels = {ins[1]: i for i, ins in enumerate(self.ins) if ins[0] == "label"}
        pc = 0
        while pc < len(self.ins):
            op, dst, srcs, guard = self.ins[pc]
            pc += 1
            if op == "label":
                continue
            if guard is not None:
                g = pred[guard.lstrip("!")]
                if guard.startswith("!"):
                    g = not g
                if not g:
                    continue
            if op == "bra":
                pc = labels[dst]
                continue
            v = [val(s) for s in srcs if not (isinstance(s, str) and s.startswith("%p"))]
            if op in ("setp.gt", "setp.eq", "setp.lt"):
                pred[dst] = {"setp.gt": v[0] > v[1], "setp.eq": v[0] == v[1], "setp.lt": v[0] < v[1]}[op]
                continue
            if op in ("mul.lo", "mul.hi"):
                p = v[0] * v[1]
                env[dst] = (p & M32) if op == "mul.lo" else (p >> 32)
                continue
            if op == "mov":
                env[dst] = v[0]
                continue
            if op == "setp.ne":
                pred[dst] = v[0] != v[1]
                continue
            if op == "selp":
                env[dst] = v[0] if pred[srcs[2]] else v[1]
                continue
            if op == "and":
                env[dst] = v[0] & v[1]
                continue
            if op == "shr":
                env[dst] = v[0] >> v[1]
                continue
            base, *mods = op.split(".")
            if base in ("mad", "madc"):
                half = mods[0]
                p = v[0] * v[1]
                p = (p & M32) if half == "lo" else (p >> 32)
                s = p + v[2] + (cf if base == "madc" else 0)
                env[dst] = s & M32
                if "cc" in mods:
                    cf = s >> 32
                else:
                    assert s >> 32 == 0, "lost carry in " + op
                continue
            if base in ("add", "addc"):
                s = v[0] + v[1] + (cf if base == "addc" else 0)
                env[dst] = s & M32
                if "cc" in mods:
                    cf = s >> 32
                else:
                    assert s >> 32 == 0, "lost carry in " + op
                continue
            if base in ("sub", "subc"):
                s = v[0] - v[1] - (cf if base == "subc" else 0)
                env[dst] = s & M32
                if "cc" in mods:
                    cf = 1 if s < 0 else 0
                continue
            raise ValueError(op)
        return env

    # -- PTX text ---------------------------------------------------------------------------
    def ptx(self):
        opmap = {}
        for i, r in enumerate(self.outputs + self.inputs):
            opmap[r] = "%%%d" % i
        lines = []
        if self.ntmp:
            lines.append(".reg .u32 t<%d>;" % self.ntmp)
        npred = len({d for op, d, _, _ in self.ins if op.startswith("setp")})
        if npred:
            lines.append(".reg .pred p<%d>;" % npred)

        def o(s):
            if isinstance(s, int):
                return "0x%08x" % s if s > 9 else str(s)
            if s.startswith("%p"):
                return s[1:]
            return opmap.get(s, s)

        for op, dst, srcs, guard in self.ins:
            g = ""
            if guard is not None:
                g = "@%s%s " % ("!" if guard.startswith("!") else "", guard.lstrip("!")[1:])
            if op == "label":
                lines.append("%s:" % dst)
            elif op == "bra":
                lines.append("%sbra %s;" % (g, dst))
            elif op.startswith("setp"):
                lines.append("%s%s.u32 %s, %s;" % (g, op, o(dst), ", ".join(o(s) for s in srcs)))
            elif op == "selp":
                lines.append("%sselp.u32 %s, %s;" % (g, o(dst), ", ".join(o(s) for s in srcs)))
            elif op == "and":
                lines.append("%sand.b32 %s, %s;" % (g, o(dst), ", ".join(o(s) for s in srcs)))
            elif op == "shr":
                lines.append("%sshr.u32 %s, %s;" % (g, o(dst), ", ".join(o(s) for s in srcs)))
            else:
                lines.append("%s%s.u32 %s, %s;" % (g, op, o(dst), ", ".join(o(s) for s in srcs)))
        return lines


class Acc:
    """One parity accumulator: limb position -> register holding it (absent = zero)."""

    def __init__(self):
        self.r = {}


def chain(P, acc, pos, mults, fixed, carry_in, top_cannot_overflow=False):
    """Add sum_j mults[j]*fixed*2^(32*(pos+2j)) into acc limbs pos..pos+2*len-1 as ONE carry chain.

    mults: list of registers/immediates (a_j or q_j); fixed: register (b_i or k).
    Returns True if a carry-out is pending in CF after the chain (the caller must consume it with
    the very next instruction).  A fresh top limb (hi half + carry <= 2^32-1) never carries out;
    top_cannot_overflow asserts the same for an existing top limb (magnitude argument at the call
    site, re-checked by the simulator, which traps on any lost carry).
    """
    n = len(mults)
    top = pos + 2 * n - 1
    if not carry_in and all((pos + t) not in acc.r for t in range(2 * n)):
        for j in range(n):                     # nothing to add to: plain products, no carries
            # written as mad.lo.cc / madc.hi with a zero addend: ptxas fuses that pair into ONE IMAD.WIDE.U32 (4.1 clk),
            # whereas mul.lo + mul.hi stay IMAD + IMAD.HI (2.0 + 4.6 clk)
            lo, hi = P.tmp(), P.tmp()
            P.emit("mad.lo.cc", lo, mults[j], fixed, 0)
            P.emit("madc.hi", hi, mults[j], fixed, 0)
            acc.r[pos + 2 * j] = lo
            acc.r[pos + 2 * j + 1] = hi
        return False
    pending = not (top not in acc.r or top_cannot_overflow)
    for j in range(n):
        for t, half in enumerate(("lo", "hi")):
            p = pos + 2 * j + t
            first = (j == 0 and t == 0)
            addend = acc.r.get(p, 0)
            dst = addend if p in acc.r else P.tmp()
            base = "madc" if (carry_in or not first) else "mad"
            cc = pending or p != top
            P.emit("%s.%s%s" % (base, half, ".cc" if cc else ""), dst, mults[j], fixed, addend)
            acc.r[p] = dst
    return pending


def cond_sub_q(P, r):
    """r (8 regs, value < 2q) -> r mod q, in place: t = r - q; keep t unless it borrowed."""
    # The top limb decides except when it equals q's top limb (probability 2^-32): only then is the full
    # comparison run, behind a branch.  The subtraction itself is 8 predicated instructions, no selects.
    P.emit("setp.gt", "%p0", r[7], QL[7])      # certainly >= q
    P.emit("setp.eq", "%p1", r[7], QL[7])      # tie: compare the low 224 bits
    P.bra("L_nt", guard="!%p1")
    t = P.tmp()
    for i in range(7):
        P.emit(("sub.cc" if i == 0 else "subc.cc"), t, r[i], QL[i])
    P.emit("subc", t, 0, 0)                    # 0xffffffff if low part < q's low part
    P.emit("setp.eq", "%p0", t, 0)
    P.label("L_nt")
    for i in range(8):
        P.emit(("sub.cc" if i == 0 else ("subc.cc" if i < 7 else "subc")), r[i], r[i], QL[i], guard="%p0")


def gen_mul():
    P = Prog("fq_mul_ptx")
    a = ["a%d" % i for i in range(8)]
    b = ["b%d" % i for i in range(8)]
    r = ["r%d" % i for i in range(8)]
    P.inputs = a + b
    P.outputs = r
    accs = [Acc(), Acc()]                      # [even positions pairs, odd position pairs]
    a_even, a_odd = a[0::2], a[1::2]
    q_even, q_odd = QL[0::2], QL[1::2]
    for i in range(8):
        A, B = accs[i & 1], accs[(i + 1) & 1]
        # 1. fold B's leftover limb at position i into A[i]; the carry enters B's chain at i+1
        cin = False
        if i in B.r:
            P.emit("add.cc", A.r[i], A.r[i], B.r.pop(i))
            cin = True
        # 2. multiplication row: a_odd*b_i into B at i+1.., a_even*b_i into A at i..
        pend = chain(P, B, i + 1, a_odd, b[i], cin)
        assert not pend
        pend = chain(P, A, i, a_even, b[i], False)
        if pend:
            P.emit("addc", B.r[i + 8], B.r[i + 8], 0)
        # 3. reduction row
        k = P.tmp()
        P.emit("mul.lo", k, A.r[i], INV32)
        # B[i+8] = hi(a7*b_i)+carries < 0x30644e75 and hi(q7*k) < 0x30644e72: the top limb cannot carry out
        pend = chain(P, B, i + 1, q_odd, k, False, top_cannot_overflow=True)
        assert not pend
        pend = chain(P, A, i, q_even, k, False)
        assert pend
        P.emit("addc", B.r[i + 8], B.r[i + 8], 0)
        A.r.pop(i)                             # now zero
    # result limbs at positions 8..15
    E, O = accs[0], accs[1]
    for p in range(8, 16):
        x, y = E.r.get(p), O.r.get(p)
        op = "add.cc" if p == 8 else ("addc.cc" if p < 15 else "addc")
        P.emit(op, r[p - 8], x if x is not None else 0, y if y is not None else 0)
    cond_sub_q(P, r)
    return P


def gen_mulwide():
    P = Prog("fq_mulwide_ptx")
    a = ["a%d" % i for i in range(8)]
    b = ["b%d" % i for i in range(8)]
    t = ["r%d" % i for i in range(16)]
    P.inputs = a + b
    P.outputs = t
    accs = [Acc(), Acc()]
    a_even, a_odd = a[0::2], a[1::2]
    for i in range(8):
        A, B = accs[i & 1], accs[(i + 1) & 1]
        for acc, pos, mults in ((B, i + 1, a_odd), (A, i, a_even)):
            pend = chain(P, acc, pos, mults, b[i], False)
            if pend and pos + 8 < 16:
                c = P.tmp()
                P.emit("addc", c, 0, 0)
                acc.r[pos + 8] = c
    E, O = accs
    P.emit("mov", t[0], E.r[0])
    for p in range(1, 16):
        op = "add.cc" if p == 1 else ("addc.cc" if p < 15 else "addc")
        P.emit(op, t[p], E.r.get(p, 0), O.r.get(p, 0))
    return P


def gen_redc():
    P = Prog("fq_redc_ptx")
    t = ["a%d" % i for i in range(16)]
    r = ["r%d" % i for i in range(8)]
    P.inputs = t
    P.outputs = r
    accs = [Acc(), Acc()]
    for i in range(8):                         # low half seeds the even accumulator (copied: inputs stay read-only)
        x = P.tmp()
        P.emit("mov", x, t[i])
        accs[0].r[i] = x
    q_even, q_odd = QL[0::2], QL[1::2]
    for i in range(8):
        A, B = accs[i & 1], accs[(i + 1) & 1]
        if i in B.r:
            P.emit("add.cc", A.r[i], A.r[i], B.r.pop(i))
            cin = True
        else:
            cin = False
        k = P.tmp()
        P.emit("mul.lo", k, A.r[i], INV32)
        pend = chain(P, B, i + 1, q_odd, k, cin)
        assert not pend
        pend = chain(P, A, i, q_even, k, False)
        assert pend
        P.emit("addc", B.r[i + 8], B.r[i + 8], 0)
        A.r.pop(i)
    E, O = accs
    for p in range(8, 16):                     # m = E + O  (value <= q)
        op = "add.cc" if p == 8 else ("addc.cc" if p < 15 else "addc")
        P.emit(op, r[p - 8], E.r.get(p, 0), O.r.get(p, 0))
    for p in range(8):                         # + high half of t  (total < 2q)
        op = "add.cc" if p == 0 else ("addc.cc" if p < 7 else "addc")
        P.emit(op, r[p], r[p], t[8 + p])
    cond_sub_q(P, r)
    return P


def gen_add():
    P = Prog("fq_add_ptx")
    a = ["a%d" % i for i in range(8)]
    b = ["b%d" % i for i in range(8)]
    r = ["r%d" % i for i in range(8)]
    P.inputs = a + b
    P.outputs = r
    for i in range(8):
        P.emit("add.cc" if i == 0 else ("addc.cc" if i < 7 else "addc"), r[i], a[i], b[i])
    cond_sub_q(P, r)
    return P


def gen_sub():
    P = Prog("fq_sub_ptx")
    a = ["a%d" % i for i in range(8)]
    b = ["b%d" % i for i in range(8)]
    r = ["r%d" % i for i in range(8)]
    P.inputs = a + b
    P.outputs = r
    for i in range(8):
        P.emit("sub.cc" if i == 0 else "subc.cc", r[i], a[i], b[i])
    m = P.tmp()
    P.emit("subc", m, 0, 0)                    # all ones on borrow
    P.emit("setp.ne", "%p0", m, 0)
    for i in range(8):                         # the final carry is the intended wrap of r - b + q past 2^256
        P.emit("add.cc" if i == 0 else "addc.cc", r[i], r[i], QL[i], guard="%p0")
    return P


NQL = [((1 << 256) - Q >> (32 * i)) & M32 for i in range(8)]     # 2^256 - q
QH242 = Q >> 242                                                   # 3097
KMUL = (1 << 32) // (QH242 + 1)                                    # floor(2^32 / 3098)


def gen_lin9(minus):
    """r = 9a + b (minus=False) or 9a - b (minus=True) mod q, fully reduced: the two halves of a multiplication
    by xi = 9 + u in Fq2 (src/fields/fq2.rs:55-57), replacing four doublings/additions and a final add/sub, each
    with its own conditional subtraction.

    t = 9a + b' (b' = b or q - b) < 10q as 9 limbs: the products 9*a_j of the even limbs run as one carry chain on
    top of b', those of the odd limbs are carry-free, one 8-limb addition joins them.  Quotient estimate from
    the top 16 bits, k = ((t >> 242) * floor(2^32 / 3098)) >> 32, never above floor(t / q) and at most one below
    (t / q - k < 1.01, argued in DESIGN.md F10 and swept by the simulator), so t - k*q < 2q fits 256 bits: it is
    formed modulo 2^256 as t + k * (2^256 - q) with two more chains, then the usual conditional subtraction."""
    P = Prog("fq_9a_minus_b_ptx" if minus else "fq_9a_plus_b_ptx")
    a = ["a%d" % i for i in range(8)]
    b = ["b%d" % i for i in range(8)]
    r = ["r%d" % i for i in range(8)]
    P.inputs = a + b
    P.outputs = r
    e = [P.tmp() for _ in range(9)]
    if minus:                                  # e = q - b  (b < q: no borrow out)
        for i in range(8):
            P.emit("sub.cc" if i == 0 else ("subc.cc" if i < 7 else "subc"), e[i], QL[i], b[i])
    # even chain: e += 9 * (a0 + a2 2^64 + a4 2^128 + a6 2^192)
    for j in range(4):
        src_lo = e[2 * j] if minus else b[2 * j]
        src_hi = e[2 * j + 1] if minus else b[2 * j + 1]
        P.emit("mad.lo.cc" if j == 0 else "madc.lo.cc", e[2 * j], a[2 * j], 9, src_lo)
        P.emit("madc.hi.cc", e[2 * j + 1], a[2 * j], 9, src_hi)
    P.emit("addc", e[8], 0, 0)
    # odd products: limbs (1,2), (3,4), (5,6), (7,8), no carries between them
    o = {}
    for j in range(4):
        lo, hi = P.tmp(), P.tmp()
        P.emit("mul.lo", lo, a[2 * j + 1], 9)
        P.emit("mul.hi", hi, a[2 * j + 1], 9)
        o[2 * j + 1], o[2 * j + 2] = lo, hi
    t = [e[0]] + [P.tmp() for _ in range(8)]
    for p in range(1, 9):
        P.emit("add.cc" if p == 1 else ("addc.cc" if p < 8 else "addc"), t[p], e[p], o[p])
    # quotient estimate
    x, th, k = P.tmp(), P.tmp(), P.tmp()
    P.emit("shr", x, t[7], 18)
    P.emit("mad.lo", th, t[8], 1 << 14, x)
    P.emit("mul.hi", k, th, KMUL)
    # r = t + k * (2^256 - q)  mod 2^256
    for j in range(4):                         # even limbs of 2^256 - q
        P.emit("mad.lo.cc" if j == 0 else "madc.lo.cc", r[2 * j], NQL[2 * j], k, t[2 * j])
        P.emit("madc.hi.cc", r[2 * j + 1], NQL[2 * j], k, t[2 * j + 1])
    for j in range(4):                         # odd limbs; the product at limb 7 keeps only its low half
        P.emit("mad.lo.cc" if j == 0 else "madc.lo.cc", r[2 * j + 1], NQL[2 * j + 1], k, r[2 * j + 1])
        if j < 3:
            P.emit("madc.hi.cc", r[2 * j + 2], NQL[2 * j + 1], k, r[2 * j + 2])
    cond_sub_q(P, r)
    return P


def gen_red9():
    """r = t mod q for a 9-limb t < 160 q: the reduction tail of gen_lin9 on its own, for sums accumulated WITHOUT a conditional
    subtraction per term (the quad-lane executor, csrc/bn_coop4.cuh: up to 8 signed terms and one multiplication by xi = 9 + u,
    offset by a multiple of q so that the total is non-negative).  Same quotient estimate from the top bits,
    k = ((t >> 242) * floor(2^32 / 3098)) >> 32 with k <= floor(t / q) <= k + 1 (swept below over every top word up to 160 q)."""
    P = Prog("fq_red9_ptx")
    t = ["a%d" % i for i in range(9)]
    r = ["r%d" % i for i in range(8)]
    P.inputs = t
    P.outputs = r
    x, th, k = P.tmp(), P.tmp(), P.tmp()
    P.emit("shr", x, t[7], 18)
    P.emit("mad.lo", th, t[8], 1 << 14, x)
    P.emit("mul.hi", k, th, KMUL)
    for j in range(4):
        P.emit("mad.lo.cc" if j == 0 else "madc.lo.cc", r[2 * j], NQL[2 * j], k, t[2 * j])
        P.emit("madc.hi.cc", r[2 * j + 1], NQL[2 * j], k, t[2 * j + 1])
    for j in range(4):
        P.emit("mad.lo.cc" if j == 0 else "madc.lo.cc", r[2 * j + 1], NQL[2 * j + 1], k, r[2 * j + 1])
        if j < 3:
            P.emit("madc.hi.cc", r[2 * j + 2], NQL[2 * j + 1], k, r[2 * j + 2])
    cond_sub_q(P, r)
    return P


RED9_MAX = 160         # fq_red9_ptx accepts t < RED9_MAX * q


def gen_wide_add():
    P = Prog("wide_add_ptx")
    a = ["a%d" % i for i in range(16)]
    b = ["b%d" % i for i in range(16)]
    P.inputs = b
    P.outputs = a                              # in/out ("+r")
    for i in range(16):
        P.emit("add.cc" if i == 0 else ("addc.cc" if i < 15 else "addc"), a[i], a[i], b[i])
    return P


def gen_wide_sub():
    """a -= b over 512 bits; on borrow add q*2^256 (keeps the value in [0, q*R) when a,b < q*R)."""
    P = Prog("wide_sub_ptx")
    a = ["a%d" % i for i in range(16)]
    b = ["b%d" % i for i in range(16)]
    P.inputs = b
    P.outputs = a
    for i in range(16):
        P.emit("sub.cc" if i == 0 else "subc.cc", a[i], a[i], b[i])
    m = P.tmp()
    P.emit("subc", m, 0, 0)
    P.emit("setp.ne", "%p0", m, 0)
    for i in range(8):                         # final carry = intended wrap past 2^512
        P.emit("add.cc" if i == 0 else "addc.cc", a[8 + i], a[8 + i], QL[i], guard="%p0")
    return P


def gen_wide_sub_nb():
    """a -= b over 512 bits for a >= b (the two Karatsuba corrections of an Fq2 product never borrow): no add-back code."""
    P = Prog("wide_sub_nb_ptx")
    a = ["a%d" % i for i in range(16)]
    b = ["b%d" % i for i in range(16)]
    P.inputs = b
    P.outputs = a
    for i in range(16):
        P.emit("sub.cc" if i == 0 else ("subc.cc" if i < 15 else "subc"), a[i], a[i], b[i])
    return P


# ------------------------------------------------------------------------------------------------
def limbs(x, n):
    return [(x >> (32 * i)) & M32 for i in range(n)]


def unl(env, names):
    return sum(env[n] << (32 * i) for i, n in enumerate(names))


def check():
    rng = random.Random(254)
    R = 1 << 256
    Rinv = pow(R, -1, Q)
    edge = [0, 1, Q - 1, Q - 2, (1 << 253), R % Q, M32, Q >> 1]
    vals = edge + [rng.randrange(Q) for _ in range(300)]
    pm, pw, pr, pa, ps = gen_mul(), gen_mulwide(), gen_redc(), gen_add(), gen_sub()
    pwa, pws, pwn = gen_wide_add(), gen_wide_sub(), gen_wide_sub_nb()
    n = 0
    for a in vals:
        for b in (vals if a in edge else rng.sample(vals, 12)):
            env = {("a%d" % i): v for i, v in enumerate(limbs(a, 8))}
            env.update({("b%d" % i): v for i, v in enumerate(limbs(b, 8))})
            assert unl(pm.run(env), pm.outputs) == (a * b * Rinv) % Q, ("mul", hex(a), hex(b))
            assert unl(pw.run(env), pw.outputs) == a * b, ("mulwide", hex(a), hex(b))
            assert unl(pa.run(env), pa.outputs) == (a + b) % Q
            assert unl(ps.run(env), ps.outputs) == (a - b) % Q
            n += 1
    # tie cases of the conditional subtraction: a + b has the same top limb as q
    top = QL[7] << 224
    for ssum in [Q - 2, Q - 1, Q, Q + 1, Q + 2, top, top + 1, Q + (1 << 100), Q - (1 << 100), top + (Q & ((1 << 224) - 1)) - 1,
                 (top | ((1 << 224) - 1))] + [top + rng.randrange(1 << 224) for _ in range(200)]:
        for _ in range(4):
            lo = max(0, ssum - (Q - 1))
            a = rng.randrange(lo, min(ssum, Q - 1) + 1)
            b = ssum - a
            assert 0 <= a < Q and 0 <= b < Q
            env = {("a%d" % i): v for i, v in enumerate(limbs(a, 8))}
            env.update({("b%d" % i): v for i, v in enumerate(limbs(b, 8))})
            assert unl(pa.run(env), pa.outputs) == (a + b) % Q, ("add tie", hex(a), hex(b))
            n += 1
    for _ in range(2000):
        t = rng.choice([rng.randrange(Q * R), Q * R - 1, 0, rng.randrange(Q) * rng.randrange(Q), (Q - 1) * (Q - 1),
                        rng.randrange(1 << 256), (rng.randrange(Q) << 256) | ((1 << 256) - 1)])
        env = {("a%d" % i): v for i, v in enumerate(limbs(t, 16))}
        assert unl(pr.run(env), pr.outputs) == (t * Rinv) % Q, ("redc", hex(t))
        u = rng.choice([rng.randrange(Q * R), 0, Q * R - 1, t])
        env = {("a%d" % i): v for i, v in enumerate(limbs(t, 16))}
        env.update({("b%d" % i): v for i, v in enumerate(limbs(u, 16))})
        got = unl(pws.run(env), pws.outputs)
        assert got == (t - u if t >= u else t - u + Q * R), "wide_sub"
        if t + u < (1 << 512):
            assert unl(pwa.run(env), pwa.outputs) == t + u, "wide_add"
        if t >= u:
            assert unl(pwn.run(env), pwn.outputs) == t - u, "wide_sub_nb"
    pp, pn = gen_lin9(False), gen_lin9(True)
    nl = 0
    # sweep the quotient estimate: 9a + b (and 9a + q - b) just below / at / above every multiple of q, plus random pairs
    targets = []
    for kq in range(0, 11):
        for d in (-2, -1, 0, 1, 2, (1 << 224) - 1, -(1 << 224), 1 << 242, -(1 << 242)):
            targets.append(kq * Q + d)
    for tgt in targets:
        for _ in range(6):
            # 9a + b = tgt with 0 <= a, b < q
            a_hi = min(Q - 1, tgt // 9) if tgt >= 0 else -1
            a_lo = max(0, -(-(tgt - (Q - 1)) // 9))
            if a_hi < a_lo:
                continue
            a9 = rng.randrange(a_lo, a_hi + 1)
            bb = tgt - 9 * a9
            if not 0 <= bb < Q:
                continue
            env = {("a%d" % i): v for i, v in enumerate(limbs(a9, 8))}
            env.update({("b%d" % i): v for i, v in enumerate(limbs(bb, 8))})
            assert unl(pp.run(env), pp.outputs) == (9 * a9 + bb) % Q, ("9a+b", hex(a9), hex(bb))
            nb = (Q - bb) % Q if bb else 0          # 9a - nb = 9a + bb - q  (same t when bb != 0)
            env.update({("b%d" % i): v for i, v in enumerate(limbs(nb, 8))})
            assert unl(pn.run(env), pn.outputs) == (9 * a9 - nb) % Q, ("9a-b", hex(a9), hex(nb))
            nl += 2
    for a9 in vals:
        for bb in (vals if a9 in edge else rng.sample(vals, 12)):
            env = {("a%d" % i): v for i, v in enumerate(limbs(a9, 8))}
            env.update({("b%d" % i): v for i, v in enumerate(limbs(bb, 8))})
            assert unl(pp.run(env), pp.outputs) == (9 * a9 + bb) % Q, ("9a+b", hex(a9), hex(bb))
            assert unl(pn.run(env), pn.outputs) == (9 * a9 - bb) % Q, ("9a-b", hex(a9), hex(bb))
            nl += 2
    # the estimate itself, exhaustively over the 16-bit top word: k <= floor(t/q) <= k + 1 for every t with that top
    for th in range(0, (10 * Q >> 242) + 2):
        k = (th * KMUL) >> 32
        lo_t, hi_t = th << 242, min(((th + 1) << 242) - 1, 10 * Q - 1)
        if lo_t > hi_t:
            continue
        assert k <= lo_t // Q and hi_t // Q <= k + 1 and hi_t - k * Q < 2 * Q, th
    p9 = gen_red9()
    n9 = 0
    cases = []
    for kq in range(0, RED9_MAX + 1):
        for d in (-2, -1, 0, 1, 2, (1 << 224) - 1, -(1 << 224), 1 << 242, -(1 << 242)):
            cases.append(kq * Q + d)
    cases += [rng.randrange(RED9_MAX * Q) for _ in range(3000)]
    for tv in cases:
        if not 0 <= tv < RED9_MAX * Q:
            continue
        env = {("a%d" % i): v for i, v in enumerate(limbs(tv, 9))}
        assert unl(p9.run(env), p9.outputs) == tv % Q, ("red9", hex(tv))
        n9 += 1
    for th in range(0, (RED9_MAX * Q >> 242) + 2):
        k = (th * KMUL) >> 32
        lo_t, hi_t = th << 242, min(((th + 1) << 242) - 1, RED9_MAX * Q - 1)
        if lo_t > hi_t:
            continue
        assert k <= lo_t // Q and hi_t // Q <= k + 1 and hi_t - k * Q < 2 * Q, th
    print("simulated OK: %d mul/mulwide/add/sub pairs, 2000 redc/wide cases, %d 9a+-b cases, %d 9-limb reductions" % (n, nl, n9))
    for p in (pm, pw, pr, pa, ps, pp, pn, p9):
        ops = [op.split(".")[0] for op, _, _, _ in p.ins]
        print("  %-16s %3d PTX instrs (%d mad/mul)" % (p.name, len(p.ins), sum(o in ("mad", "madc", "mul") for o in ops)))


def emit_function(P, inout=False):
    nout, nin = len(P.outputs), len(P.inputs)
    lines = P.ptx()
    body = "\n".join('        "%s\\n\\t"' % ln for ln in lines)
    if inout:
        sig = "uint32_t (&a)[%d], const uint32_t (&b)[%d]" % (nout, nin)
        outs = ", ".join('"+r"(a[%d])' % i for i in range(nout))
        ins = ", ".join('"r"(b[%d])' % i for i in range(nin))
    elif P.name == "fq_red9_ptx":
        sig = "uint32_t (&r)[8], const uint32_t (&t)[9]"
        outs = ", ".join('"=&r"(r[%d])' % i for i in range(8))
        ins = ", ".join('"r"(t[%d])' % i for i in range(9))
    elif P.name == "fq_redc_ptx":
        sig = "uint32_t (&r)[8], const uint32_t (&t)[16]"
        outs = ", ".join('"=&r"(r[%d])' % i for i in range(8))
        ins = ", ".join('"r"(t[%d])' % i for i in range(16))
    else:
        sig = "uint32_t (&r)[%d], const uint32_t (&a)[8], const uint32_t (&b)[8]" % nout
        outs = ", ".join('"=&r"(r[%d])' % i for i in range(nout))
        ins = ", ".join('"r"(a[%d])' % i for i in range(8)) + ", " + ", ".join('"r"(b[%d])' % i for i in range(8))
    return ("__device__ __forceinline__ void %s(%s) {\n    asm(\"{\\n\\t\"\n%s\n        \"}\"\n        : %s\n        : %s);\n}\n"
            % (P.name, sig, body, outs, ins))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--check", action="store_true", help="simulate only")
    ap.add_argument("--out", default=None, help="write the header here instead of csrc/fq_ptx.cuh")
    args = ap.parse_args()
    check()
    if args.check:
        return 0
    out = args.out or os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "paritytech-bn_b200", "csrc", "fq_ptx.cuh")
    with open(out, "w") as f:
        f.write("// GENERATED by tools/gen_fq_ptx.py -- do not edit.  Simulated against big-int ground truth at\n"
                "// generation time.  8x32-bit-limb BN254 Fq arithmetic as single inline-PTX carry-chain blocks.\n"
                "// Outputs are early-clobber ('=&r'): they never alias an input register.\n"
                "#pragma once\n#include <stdint.h>\n\n")
        for P, io in ((gen_mul(), False), (gen_mulwide(), False), (gen_redc(), False), (gen_add(), False), (gen_sub(), False), (gen_lin9(False), False), (gen_lin9(True), False), (gen_red9(), False),
                      (gen_wide_add(), True), (gen_wide_sub(), True), (gen_wide_sub_nb(), True)):
            f.write(emit_function(P, io))
            f.write("\n")
    print("wrote", os.path.normpath(out))
    return 0


if __name__ == "__main__":
    sys.exit(main())
