"""CPU oracle (Python big-int) for the BN254 pairing path of substrate-bn.  TEST INFRASTRUCTURE ONLY.

This is a restatement of the reference crate's algorithm, formula for formula, on canonical
(non-Montgomery) Python integers.  Only tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline leg may import it; the product path (paritytech-bn_b200/) never does.

Parity status: PINNED.  tests/test_oracle_golden.py checks this module against every
known-answer vector the reference's own tests hold for the path (fixture
tests/golden/reference_vectors.json, extracted by tests/golden/extract_golden.py):
test_str, fq12_test_vector, test_cyclotomic_exp, test_prepared_g2 (all 87 EllCoeffs),
test_miller_loop, test_reduced_pairing, and the Montgomery constant literals.

Every function cites the reference file:line (relative to /root/reference) it follows.
Representation: Fq = int in [0,q); Fq2 = (c0, c1); Fq6 = (c0, c1, c2) of Fq2; Fq12 = (c0, c1) of
Fq6; Jacobian point = (x, y, z); affine point = (x, y).  Montgomery bytes (the crate's repr(C)
memory image, R = 2^256) are produced only by the *_to_bytes / *_from_bytes helpers.
"""

# ---------------------------------------------------------------------------------------------
# constants (src/fields/fp.rs:166-222, src/fields/fq12.rs:122, src/groups/mod.rs:9-14)
# ---------------------------------------------------------------------------------------------
Q = 0x30644e72e131a029b85045b68181585d97816a916871ca8d3c208c16d87cfd47
R_ORDER = 0x30644e72e131a029b85045b68181585d2833e84879b9709143e1f593f0000001
BN_U = 4965661367192848881
MONT_R = 1 << 256
ATE_LOOP_COUNT_NAF = [1, 0, 1, 0, 0, 0, 3, 0, 3, 0, 0, 0, 3, 0, 1, 0, 3, 0, 0, 3, 0, 0, 0, 0, 0, 1, 0, 0, 3, 0, 1, 0,
                      0, 3, 0, 0, 0, 0, 3, 0, 1, 0, 0, 0, 3, 0, 3, 0, 0, 1, 0, 0, 0, 3, 0, 0, 3, 0, 1, 0, 1, 0, 0, 0]

assert Q == 36 * BN_U**4 + 36 * BN_U**3 + 24 * BN_U**2 + 6 * BN_U + 1
assert R_ORDER == 36 * BN_U**4 + 36 * BN_U**3 + 18 * BN_U**2 + 6 * BN_U + 1


def _naf_value():
    v = 1
    for d in ATE_LOOP_COUNT_NAF:
        v = 2 * v + (1 if d == 1 else -1 if d == 3 else 0)
    return v


assert _naf_value() == 6 * BN_U + 2 == 0x19d797039be763ba8

# ---------------------------------------------------------------------------------------------
# Fq (src/arith.rs:281-316 modular add/sub/neg/mul; src/fields/fp.rs:108-117 inverse)
# ---------------------------------------------------------------------------------------------


def fq_inv(a):
    """Fq::inverse, src/fields/fp.rs:108-117 (None for zero).  Any exact inverse is bit-identical."""
    if a % Q == 0:
        return None
    return pow(a, -1, Q)


# ---------------------------------------------------------------------------------------------
# Fq2 = Fq[u]/(u^2+1)   (src/fields/fq2.rs)
# ---------------------------------------------------------------------------------------------
FQ2_ZERO = (0, 0)
FQ2_ONE = (1, 0)
XI = (9, 1)  # fq2_nonresidue, src/fields/fq2.rs:18-34


def fq2_add(a, b):
    return ((a[0] + b[0]) % Q, (a[1] + b[1]) % Q)


def fq2_sub(a, b):
    return ((a[0] - b[0]) % Q, (a[1] - b[1]) % Q)


def fq2_neg(a):
    return ((-a[0]) % Q, (-a[1]) % Q)


def fq2_mul(a, b):
    """src/fields/fq2.rs:133-149 (Karatsuba, beta = -1)."""
    aa = a[0] * b[0]
    bb = a[1] * b[1]
    return ((aa - bb) % Q, ((a[0] + a[1]) * (b[0] + b[1]) - aa - bb) % Q)


def fq2_sqr(a):
    """src/fields/fq2.rs:105-117 (complex squaring)."""
    ab = a[0] * a[1]
    return (((a[0] - a[1]) * (a[0] + a[1])) % Q, (ab + ab) % Q)


def fq2_scale(a, k):
    """src/fields/fq2.rs:48-53."""
    return ((a[0] * k) % Q, (a[1] * k) % Q)


def fq2_mul_by_nonresidue(a):
    """src/fields/fq2.rs:55-57: a * (9 + u)."""
    return ((9 * a[0] - a[1]) % Q, (9 * a[1] + a[0]) % Q)


def fq2_frobenius(a, power):
    """src/fields/fq2.rs:59-68: conjugation for odd powers."""
    return a if power % 2 == 0 else (a[0], (-a[1]) % Q)


def fq2_inv(a):
    """src/fields/fq2.rs:119-130."""
    t = fq_inv((a[0] * a[0] + a[1] * a[1]) % Q)
    if t is None:
        return None
    return ((a[0] * t) % Q, (-(a[1] * t)) % Q)


def fq2_is_zero(a):
    return a[0] == 0 and a[1] == 0


def fq2_pow(a, e):
    """FieldElement::pow over a 256-bit exponent, src/fields/mod.rs:35-46."""
    res = FQ2_ONE
    for i in range(255, -1, -1):
        res = fq2_sqr(res)
        if (e >> i) & 1:
            res = fq2_mul(a, res)
    return res


# ---------------------------------------------------------------------------------------------
# Fq6 = Fq2[v]/(v^3 - xi)   (src/fields/fq6.rs)
# ---------------------------------------------------------------------------------------------
FQ6_ZERO = (FQ2_ZERO, FQ2_ZERO, FQ2_ZERO)
FQ6_ONE = (FQ2_ONE, FQ2_ZERO, FQ2_ZERO)

# Frobenius coefficients, derived (not copied): gamma_{1,k} = xi^((q^k - 1)/3), gamma_{2,k} = xi^(2(q^k-1)/3)
# for Fq6 (src/fields/fq6.rs:5-90) and xi^((q^k-1)/6) for Fq12 (src/fields/fq12.rs:6-48).


def _fq2_pow_big(a, e):
    res = FQ2_ONE
    for bit in bin(e)[2:]:
        res = fq2_sqr(res)
        if bit == "1":
            res = fq2_mul(res, a)
    return res


FQ6_FROB_C1 = {k: _fq2_pow_big(XI, (Q**k - 1) // 3) for k in (0, 1, 2, 3)}
FQ6_FROB_C2 = {k: _fq2_pow_big(XI, 2 * (Q**k - 1) // 3) for k in (0, 1, 2, 3)}
FQ12_FROB_C1 = {k: _fq2_pow_big(XI, (Q**k - 1) // 6) for k in (0, 1, 2, 3)}


def fq6_add(a, b):
    return (fq2_add(a[0], b[0]), fq2_add(a[1], b[1]), fq2_add(a[2], b[2]))


def fq6_sub(a, b):
    return (fq2_sub(a[0], b[0]), fq2_sub(a[1], b[1]), fq2_sub(a[2], b[2]))


def fq6_neg(a):
    return (fq2_neg(a[0]), fq2_neg(a[1]), fq2_neg(a[2]))


def fq6_mul_by_nonresidue(a):
    """src/fields/fq6.rs:109-115: multiply by v."""
    return (fq2_mul_by_nonresidue(a[2]), a[0], a[1])


def fq6_scale(a, k):
    """src/fields/fq6.rs:117-123 (k in Fq2)."""
    return (fq2_mul(a[0], k), fq2_mul(a[1], k), fq2_mul(a[2], k))


def fq6_frobenius(a, power):
    """src/fields/fq6.rs:125-131 (powers 0..3 only, like the reference)."""
    if power not in (0, 1, 2, 3):
        raise NotImplementedError("reference has unimplemented!() here, src/fields/fq6.rs:45,88")
    return (fq2_frobenius(a[0], power),
            fq2_mul(fq2_frobenius(a[1], power), FQ6_FROB_C1[power]),
            fq2_mul(fq2_frobenius(a[2], power), FQ6_FROB_C2[power]))


def fq6_mul(a, b):
    """src/fields/fq6.rs:194-208."""
    a_a = fq2_mul(a[0], b[0])
    b_b = fq2_mul(a[1], b[1])
    c_c = fq2_mul(a[2], b[2])
    c0 = fq2_add(fq2_mul_by_nonresidue(
        fq2_sub(fq2_sub(fq2_mul(fq2_add(a[1], a[2]), fq2_add(b[1], b[2])), b_b), c_c)), a_a)
    c1 = fq2_add(fq2_sub(fq2_sub(fq2_mul(fq2_add(a[0], a[1]), fq2_add(b[0], b[1])), a_a), b_b),
                 fq2_mul_by_nonresidue(c_c))
    c2 = fq2_sub(fq2_add(fq2_sub(fq2_mul(fq2_add(a[0], a[2]), fq2_add(b[0], b[2])), a_a), b_b), c_c)
    return (c0, c1, c2)


def fq6_sqr(a):
    """src/fields/fq6.rs:163-177."""
    s0 = fq2_sqr(a[0])
    ab = fq2_mul(a[0], a[1])
    s1 = fq2_add(ab, ab)
    s2 = fq2_sqr(fq2_add(fq2_sub(a[0], a[1]), a[2]))
    bc = fq2_mul(a[1], a[2])
    s3 = fq2_add(bc, bc)
    s4 = fq2_sqr(a[2])
    return (fq2_add(s0, fq2_mul_by_nonresidue(s3)),
            fq2_add(s1, fq2_mul_by_nonresidue(s4)),
            fq2_sub(fq2_sub(fq2_add(fq2_add(s1, s2), s3), s0), s4))


def fq6_inv(a):
    """src/fields/fq6.rs:179-191."""
    c0 = fq2_sub(fq2_sqr(a[0]), fq2_mul(a[1], fq2_mul_by_nonresidue(a[2])))
    c1 = fq2_sub(fq2_mul_by_nonresidue(fq2_sqr(a[2])), fq2_mul(a[0], a[1]))
    c2 = fq2_sub(fq2_sqr(a[1]), fq2_mul(a[0], a[2]))
    t = fq2_inv(fq2_add(fq2_mul_by_nonresidue(fq2_add(fq2_mul(a[2], c1), fq2_mul(a[1], c2))),
                        fq2_mul(a[0], c0)))
    if t is None:
        return None
    return (fq2_mul(t, c0), fq2_mul(t, c1), fq2_mul(t, c2))


def fq6_is_zero(a):
    return all(fq2_is_zero(c) for c in a)


# ---------------------------------------------------------------------------------------------
# Fq12 = Fq6[w]/(w^2 - v)   (src/fields/fq12.rs)
# ---------------------------------------------------------------------------------------------
FQ12_ZERO = (FQ6_ZERO, FQ6_ZERO)
FQ12_ONE = (FQ6_ONE, FQ6_ZERO)


def fq12_add(a, b):
    return (fq6_add(a[0], b[0]), fq6_add(a[1], b[1]))


def fq12_sub(a, b):
    return (fq6_sub(a[0], b[0]), fq6_sub(a[1], b[1]))


def fq12_neg(a):
    return (fq6_neg(a[0]), fq6_neg(a[1]))


def fq12_mul(a, b):
    """src/fields/fq12.rs:316-328."""
    aa = fq6_mul(a[0], b[0])
    bb = fq6_mul(a[1], b[1])
    return (fq6_add(fq6_mul_by_nonresidue(bb), aa),
            fq6_sub(fq6_sub(fq6_mul(fq6_add(a[0], a[1]), fq6_add(b[0], b[1])), aa), bb))


def fq12_sqr(a):
    """src/fields/fq12.rs:295-303 (complex squaring)."""
    ab = fq6_mul(a[0], a[1])
    c0 = fq6_sub(fq6_sub(fq6_mul(fq6_add(fq6_mul_by_nonresidue(a[1]), a[0]), fq6_add(a[0], a[1])), ab),
                 fq6_mul_by_nonresidue(ab))
    return (c0, fq6_add(ab, ab))


def fq12_inv(a):
    """src/fields/fq12.rs:305-313."""
    t = fq6_inv(fq6_sub(fq6_sqr(a[0]), fq6_mul_by_nonresidue(fq6_sqr(a[1]))))
    if t is None:
        return None
    return (fq6_mul(a[0], t), fq6_neg(fq6_mul(a[1], t)))


def fq12_unitary_inverse(a):
    """src/fields/fq12.rs:126-128."""
    return (a[0], fq6_neg(a[1]))


def fq12_frobenius(a, power):
    """src/fields/fq12.rs:112-119."""
    return (fq6_frobenius(a[0], power), fq6_scale(fq6_frobenius(a[1], power), FQ12_FROB_C1[power]))


def fq12_mul_by_024(f, ell_0, ell_vw, ell_vv):
    """src/fields/fq12.rs:130-196.  Sparse multiply, statement for statement."""
    z0, z1, z2 = f[0]
    z3, z4, z5 = f[1]
    x0, x2, x4 = ell_0, ell_vv, ell_vw
    A, S, M, NR = fq2_add, fq2_sub, fq2_mul, fq2_mul_by_nonresidue

    d0 = M(z0, x0)
    d2 = M(z2, x2)
    d4 = M(z4, x4)
    t2 = A(z0, z4)
    t1 = A(z0, z2)
    s0 = A(A(z1, z3), z5)

    s1 = M(z1, x2)
    t3 = A(s1, d4)
    t4 = A(NR(t3), d0)
    z0 = t4

    t3 = M(z5, x4)
    s1 = A(s1, t3)
    t3 = A(t3, d2)
    t4 = NR(t3)
    t3 = M(z1, x0)
    s1 = A(s1, t3)
    t4 = A(t4, t3)
    z1 = t4

    t0 = A(x0, x2)
    t3 = S(S(M(t1, t0), d0), d2)
    t4 = M(z3, x4)
    s1 = A(s1, t4)
    t3 = A(t3, t4)

    t0 = A(z2, z4)
    z2 = t3

    t1 = A(x2, x4)
    t3 = S(S(M(t0, t1), d2), d4)
    t4 = NR(t3)
    t3 = M(z3, x0)
    s1 = A(s1, t3)
    t4 = A(t4, t3)
    z3 = t4

    t3 = M(z5, x2)
    s1 = A(s1, t3)
    t4 = NR(t3)
    t0 = A(x0, x4)
    t3 = S(S(M(t2, t0), d0), d4)
    t4 = A(t4, t3)
    z4 = t4

    t0 = A(A(x0, x2), x4)
    t3 = S(M(s0, t0), s1)
    z5 = t3
    return ((z0, z1, z2), (z3, z4, z5))


def fq12_cyclotomic_squared(a):
    """src/fields/fq12.rs:198-247 (Granger-Scott; note the coefficient relabelling)."""
    z0, z4, z3 = a[0]
    z2, z1, z5 = a[1]
    A, S, M, NR = fq2_add, fq2_sub, fq2_mul, fq2_mul_by_nonresidue

    tmp = M(z0, z1)
    t0 = S(S(M(A(z0, z1), A(NR(z1), z0)), tmp), NR(tmp))
    t1 = A(tmp, tmp)
    tmp = M(z2, z3)
    t2 = S(S(M(A(z2, z3), A(NR(z3), z2)), tmp), NR(tmp))
    t3 = A(tmp, tmp)
    tmp = M(z4, z5)
    t4 = S(S(M(A(z4, z5), A(NR(z5), z4)), tmp), NR(tmp))
    t5 = A(tmp, tmp)

    z0 = S(t0, z0); z0 = A(z0, z0); z0 = A(z0, t0)
    z1 = A(t1, z1); z1 = A(z1, z1); z1 = A(z1, t1)
    tmp = NR(t5)
    z2 = A(tmp, z2); z2 = A(z2, z2); z2 = A(z2, tmp)
    z3 = S(t4, z3); z3 = A(z3, z3); z3 = A(z3, t4)
    z4 = S(t2, z4); z4 = A(z4, z4); z4 = A(z4, t2)
    z5 = A(t3, z5); z5 = A(z5, z5); z5 = A(z5, t3)
    return ((z0, z4, z3), (z2, z1, z5))


def fq12_cyclotomic_pow(a, e):
    """src/fields/fq12.rs:249-266 (MSB first over 256 bits, skipping leading zeros)."""
    res = FQ12_ONE
    found_one = False
    for i in range(255, -1, -1):
        if found_one:
            res = fq12_cyclotomic_squared(res)
        if (e >> i) & 1:
            found_one = True
            res = fq12_mul(a, res)
    return res


def fq12_exp_by_neg_z(a):
    """src/fields/fq12.rs:121-124."""
    return fq12_unitary_inverse(fq12_cyclotomic_pow(a, BN_U))


def fq12_final_exp_first_chunk(a):
    """src/fields/fq12.rs:62-73."""
    b = fq12_inv(a)
    if b is None:
        return None
    c = fq12_mul(fq12_unitary_inverse(a), b)
    return fq12_mul(fq12_frobenius(c, 2), c)


def fq12_final_exp_last_chunk(x):
    """src/fields/fq12.rs:75-105."""
    a = fq12_exp_by_neg_z(x)
    b = fq12_cyclotomic_squared(a)
    c = fq12_cyclotomic_squared(b)
    d = fq12_mul(c, b)
    e = fq12_exp_by_neg_z(d)
    f = fq12_cyclotomic_squared(e)
    g = fq12_exp_by_neg_z(f)
    h = fq12_unitary_inverse(d)
    i = fq12_unitary_inverse(g)
    j = fq12_mul(i, e)
    k = fq12_mul(j, h)
    l = fq12_mul(k, b)
    m = fq12_mul(k, e)
    n = fq12_mul(x, m)
    o = fq12_frobenius(l, 1)
    p = fq12_mul(o, n)
    q = fq12_frobenius(k, 2)
    r = fq12_mul(q, p)
    s = fq12_unitary_inverse(x)
    t = fq12_mul(s, l)
    u = fq12_frobenius(t, 3)
    return fq12_mul(u, r)


def fq12_final_exponentiation(a):
    """src/fields/fq12.rs:107-110 (None iff a == 0)."""
    fc = fq12_final_exp_first_chunk(a)
    return None if fc is None else fq12_final_exp_last_chunk(fc)


def fq12_pow(a, e):
    """FieldElement::pow, src/fields/mod.rs:35-46 (256 squarings, no leading-zero skip) -> Gt::pow."""
    res = FQ12_ONE
    for i in range(255, -1, -1):
        res = fq12_sqr(res)
        if (e >> i) & 1:
            res = fq12_mul(a, res)
    return res


# ---------------------------------------------------------------------------------------------
# Groups (src/groups/mod.rs:198-369).  Generic over the base field through a tiny op table.
# ---------------------------------------------------------------------------------------------


class _Ops:
    def __init__(self, add, sub, neg, mul, sqr, inv, zero, one, is_zero):
        self.add, self.sub, self.neg, self.mul, self.sqr, self.inv = add, sub, neg, mul, sqr, inv
        self.zero, self.one, self.is_zero = zero, one, is_zero


FQ_OPS = _Ops(lambda a, b: (a + b) % Q, lambda a, b: (a - b) % Q, lambda a: (-a) % Q,
              lambda a, b: (a * b) % Q, lambda a: (a * a) % Q, fq_inv, 0, 1, lambda a: a == 0)
FQ2_OPS = _Ops(fq2_add, fq2_sub, fq2_neg, fq2_mul, fq2_sqr, fq2_inv, FQ2_ZERO, FQ2_ONE, fq2_is_zero)

G1_ONE = (1, 2, 1)  # src/groups/mod.rs:381-392
G1_B = 3            # src/groups/mod.rs:394-401
# G2 generator: canonical values of the Montgomery literals at src/groups/mod.rs:418-449; the same
# numbers appear in decimal in the reference's predefined_pair test (:1010-1021).
G2_ONE = ((10857046999023057135944570762232829481370756359578518086990519993285655852781,
           11559732032986387107991004021392285783925812861821192530917403151452391805634),
          (8495653923123431417604973247489272438418190587263600148770280649306958101930,
           4082367875863433681332203403145435568316851327593401208105741076214120093531),
          FQ2_ONE)
G2_B = fq2_mul((3, 0), fq2_inv(XI))  # b' = 3/xi, src/groups/mod.rs:452-467
TWO_INV = fq_inv(2)                  # src/groups/mod.rs:520-528
TWIST_MUL_BY_Q_X = _fq2_pow_big(XI, (Q - 1) // 3)  # src/groups/mod.rs:530-546
TWIST_MUL_BY_Q_Y = _fq2_pow_big(XI, (Q - 1) // 2)  # src/groups/mod.rs:548-564


def g_zero(F):
    """src/groups/mod.rs:230-236: (0, 1, 0)."""
    return (F.zero, F.one, F.zero)


def g_is_zero(F, p):
    return F.is_zero(p[2])


def g_double(F, p):
    """src/groups/mod.rs:250-269 (dbl-2009-l)."""
    x, y, z = p
    a = F.sqr(x)
    b = F.sqr(y)
    c = F.sqr(b)
    d = F.sub(F.sub(F.sqr(F.add(x, b)), a), c)
    d = F.add(d, d)
    e = F.add(F.add(a, a), a)
    f = F.sqr(e)
    x3 = F.sub(f, F.add(d, d))
    eight_c = F.add(c, c)
    eight_c = F.add(eight_c, eight_c)
    eight_c = F.add(eight_c, eight_c)
    y1z1 = F.mul(y, z)
    return (x3, F.sub(F.mul(e, F.sub(d, x3)), eight_c), F.add(y1z1, y1z1))


def g_add(F, p, o):
    """src/groups/mod.rs:294-334 (add-2007-bl with the zero / equal-point branches)."""
    if g_is_zero(F, p):
        return o
    if g_is_zero(F, o):
        return p
    z1_squared = F.sqr(p[2])
    z2_squared = F.sqr(o[2])
    u1 = F.mul(p[0], z2_squared)
    u2 = F.mul(o[0], z1_squared)
    z1_cubed = F.mul(p[2], z1_squared)
    z2_cubed = F.mul(o[2], z2_squared)
    s1 = F.mul(p[1], z2_cubed)
    s2 = F.mul(o[1], z1_cubed)
    if u1 == u2 and s1 == s2:
        return g_double(F, p)
    h = F.sub(u2, u1)
    s2_minus_s1 = F.sub(s2, s1)
    i = F.sqr(F.add(h, h))
    j = F.mul(h, i)
    r = F.add(s2_minus_s1, s2_minus_s1)
    v = F.mul(u1, i)
    s1_j = F.mul(s1, j)
    x3 = F.sub(F.sub(F.sqr(r), j), F.add(v, v))
    return (x3,
            F.sub(F.mul(r, F.sub(v, x3)), F.add(s1_j, s1_j)),
            F.mul(F.sub(F.sub(F.sqr(F.add(p[2], o[2])), z1_squared), z2_squared), h))


def g_neg(F, p):
    """src/groups/mod.rs:336-350."""
    return p if g_is_zero(F, p) else (p[0], F.neg(p[1]), p[2])


def g_mul(F, p, k):
    """src/groups/mod.rs:272-292.  k is the CANONICAL scalar (what U256::from(Fr) yields, :279)."""
    res = g_zero(F)
    found_one = False
    for i in range(255, -1, -1):
        if found_one:
            res = g_double(F, res)
        if (k >> i) & 1:
            found_one = True
            res = g_add(F, res, p)
    return res


def g_to_affine(F, p):
    """src/groups/mod.rs:198-217 (None at infinity; copy if z is exactly one)."""
    if F.is_zero(p[2]):
        return None
    if p[2] == F.one:
        return (p[0], p[1])
    zinv = F.inv(p[2])
    zinv_squared = F.sqr(zinv)
    return (F.mul(p[0], zinv_squared), F.mul(p[1], F.mul(zinv_squared, zinv)))


def g_eq(F, p, o):
    """src/groups/mod.rs:169-195 (projective equality)."""
    if g_is_zero(F, p):
        return g_is_zero(F, o)
    if g_is_zero(F, o):
        return False
    z1s, z2s = F.sqr(p[2]), F.sqr(o[2])
    if F.mul(p[0], z2s) != F.mul(o[0], z1s):
        return False
    return F.mul(p[1], F.mul(o[2], z2s)) == F.mul(o[1], F.mul(p[2], z1s))


def g_normalize(F, p):
    """Group::normalize, src/lib.rs:391-398 / :542-549."""
    a = g_to_affine(F, p)
    return p if a is None else (a[0], a[1], F.one)


def affine_on_curve(F, b, x, y):
    """AffineG::new on-curve test, src/groups/mod.rs:94-96."""
    return F.sqr(y) == F.add(F.mul(F.sqr(x), x), b)


def g2_in_subgroup(x, y):
    """AffineG::new order check for G2, src/groups/mod.rs:97-107: p*(-1) + p == 0."""
    p = (x, y, FQ2_ONE)
    return g_is_zero(FQ2_OPS, g_add(FQ2_OPS, g_mul(FQ2_OPS, p, R_ORDER - 1), p))


# ---------------------------------------------------------------------------------------------
# Square roots and the compressed wire formats
# ---------------------------------------------------------------------------------------------
FQ_MINUS3_DIV4 = (Q - 3) // 4   # (-3) * 4^-1 mod q, src/fields/fp.rs:233-236 (q = 3 mod 4, so this is exact)
FQ_MINUS1_DIV2 = (Q - 1) // 2   # (-1) * 2^-1 mod q, src/fields/fp.rs:238-241


def fq_sqrt(a):
    """Fq::sqrt, src/fields/fp.rs:245-260: a1 = a^((q-3)/4); None iff a1 * (a1 * a) == -1; else a1 * a."""
    a1 = pow(a, FQ_MINUS3_DIV4, Q)
    a1a = a1 * a % Q
    if a1 * a1a % Q == Q - 1:
        return None
    return a1a


def fq2_sqrt(a):
    """Fq2::sqrt, src/fields/fq2.rs:208-224 (the q = 3 mod 4 'complex method')."""
    a1 = fq2_pow(a, FQ_MINUS3_DIV4)
    a1a = fq2_mul(a1, a)
    alpha = fq2_mul(a1, a1a)
    a0 = fq2_mul(fq2_pow(alpha, Q), alpha)
    minus_one = (Q - 1, 0)
    if a0 == minus_one:
        return None
    if alpha == minus_one:
        return fq2_mul((0, 1), a1a)
    b = fq2_pow(fq2_add(alpha, FQ2_ONE), FQ_MINUS1_DIV2)
    return fq2_mul(b, a1a)


# status codes shared with include/bn254_b200.h (BN254_DECODE_*)
DECODE_OK, DECODE_INVALID_ENCODING, DECODE_NOT_MEMBER, DECODE_FIELD_NOT_MEMBER = 0, 1, 2, 3


def g1_from_compressed(b):
    """G1::from_compressed, src/lib.rs:359-375 -> (status, (x, y, 1) | None).  33 bytes: tag 2/3, x big-endian.
    The order of the checks is the reference's: field membership, square root, tag, AffineG1::new."""
    if len(b) != 33:
        return DECODE_INVALID_ENCODING, None
    sign, x = b[0], int.from_bytes(b[1:], "big")
    if x >= Q:                                   # Fq::from_slice -> CurveError::Field(NotMember), :154-159
        return DECODE_FIELD_NOT_MEMBER, None
    y = fq_sqrt((x * x * x + G1_B) % Q)
    if y is None:
        return DECODE_NOT_MEMBER, None
    if sign == 2:
        if y & 1:
            y = (-y) % Q
    elif sign == 3:
        if not y & 1:
            y = (-y) % Q
    else:
        return DECODE_INVALID_ENCODING, None
    if not affine_on_curve(FQ_OPS, G1_B, x, y):
        return DECODE_NOT_MEMBER, None
    return DECODE_OK, (x, y, 1)


def g2_from_compressed(b):
    """G2::from_compressed, src/lib.rs:506-526.  65 bytes: tag 10/11, then x as the 512-bit big-endian
    integer c1 * q + c0 (Fq2::from_slice :260-267 via U512::divrem, src/arith.rs:116-138).  y is the root
    with the smaller (tag 10) or larger (tag 11) value of c1 * q + c0 (to_u512, src/fields/fq2.rs:226-231)."""
    if len(b) != 65:
        return DECODE_INVALID_ENCODING, None
    sign, n = b[0], int.from_bytes(b[1:], "big")
    c1, c0 = divmod(n, Q)
    if c1 >= Q:                                  # quotient None or >= modulus -> Field(NotMember)
        return DECODE_FIELD_NOT_MEMBER, None
    x = (c0, c1)
    y = fq2_sqrt(fq2_add(fq2_mul(fq2_mul(x, x), x), G2_B))
    if y is None:
        return DECODE_NOT_MEMBER, None
    yn = fq2_neg(y)
    y_gt = y[1] * Q + y[0] > yn[1] * Q + yn[0]
    if sign == 10:
        e_y = yn if y_gt else y
    elif sign == 11:
        e_y = y if y_gt else yn
    else:
        return DECODE_INVALID_ENCODING, None
    if not affine_on_curve(FQ2_OPS, G2_B, x, e_y) or not g2_in_subgroup(x, e_y):
        return DECODE_NOT_MEMBER, None
    return DECODE_OK, (x, e_y, FQ2_ONE)


# ---------------------------------------------------------------------------------------------
# Pairing (src/groups/mod.rs:515-926)
# ---------------------------------------------------------------------------------------------


def g2_mul_by_q(q):
    """AffineG2::mul_by_q, src/groups/mod.rs:694-699."""
    return (fq2_mul(TWIST_MUL_BY_Q_X, fq2_frobenius(q[0], 1)), fq2_mul(TWIST_MUL_BY_Q_Y, fq2_frobenius(q[1], 1)))


def g2_doubling_step(r):
    """doubling_step_for_flipped_miller_loop, src/groups/mod.rs:754-776.  Returns (r', (ell_0, ell_vw, ell_vv))."""
    x, y, z = r
    A, S, M, SQ = fq2_add, fq2_sub, fq2_mul, fq2_sqr
    a = fq2_scale(M(x, y), TWO_INV)
    b = SQ(y)
    c = SQ(z)
    d = A(A(c, c), c)
    e = M(G2_B, d)
    f = A(A(e, e), e)
    g = fq2_scale(A(b, f), TWO_INV)
    h = S(SQ(A(y, z)), A(b, c))
    i = S(e, b)
    j = SQ(x)
    e_sq = SQ(e)
    r2 = (M(a, S(b, f)), S(SQ(g), A(A(e_sq, e_sq), e_sq)), M(b, h))
    return r2, (M(XI, i), fq2_neg(h), A(A(j, j), j))


def g2_mixed_addition_step(r, base):
    """mixed_addition_step_for_flipped_miller_loop, src/groups/mod.rs:731-752."""
    x, y, z = r
    A, S, M, SQ = fq2_add, fq2_sub, fq2_mul, fq2_sqr
    d = S(x, M(z, base[0]))
    e = S(y, M(z, base[1]))
    f = SQ(d)
    g = SQ(e)
    h = M(d, f)
    i = M(x, f)
    j = S(A(M(z, g), h), A(i, i))
    r2 = (M(d, j), S(M(e, S(i, j)), M(h, y)), M(z, h))
    return r2, (M(XI, S(M(e, base[0]), M(d, base[1]))), d, fq2_neg(e))  # (ell_0, ell_vw, ell_vv)


def g2_precompute(q):
    """AffineG2::precompute, src/groups/mod.rs:701-728.  q affine; returns (q, [87 x (ell_0, ell_vw, ell_vv)])."""
    r = (q[0], q[1], FQ2_ONE)
    coeffs = []
    q_neg = (q[0], fq2_neg(q[1]))
    for d in ATE_LOOP_COUNT_NAF:
        r, c = g2_doubling_step(r)
        coeffs.append(c)
        if d == 1:
            r, c = g2_mixed_addition_step(r, q)
            coeffs.append(c)
        if d == 3:
            r, c = g2_mixed_addition_step(r, q_neg)
            coeffs.append(c)
    q1 = g2_mul_by_q(q)
    q2 = g2_mul_by_q(q1)
    q2 = (q2[0], fq2_neg(q2[1]))
    r, c = g2_mixed_addition_step(r, q1)
    coeffs.append(c)
    r, c = g2_mixed_addition_step(r, q2)
    coeffs.append(c)
    return (q, coeffs)


def _line(f, c, p):
    return fq12_mul_by_024(f, c[0], fq2_scale(c[1], p[1]), fq2_scale(c[2], p[0]))


def miller_loop(precomp, p):
    """G2Precomp::miller_loop, src/groups/mod.rs:580-606.  p affine G1."""
    coeffs = precomp[1]
    f = FQ12_ONE
    idx = 0
    for d in ATE_LOOP_COUNT_NAF:
        f = _line(fq12_sqr(f), coeffs[idx], p)
        idx += 1
        if d != 0:
            f = _line(f, coeffs[idx], p)
            idx += 1
    f = _line(f, coeffs[idx], p)
    idx += 1
    f = _line(f, coeffs[idx], p)
    return f


def miller_loop_batch(precomps, ps):
    """groups::miller_loop_batch, src/groups/mod.rs:609-640 (zip truncates to the shorter list)."""
    pairs = list(zip(precomps, ps))
    f = FQ12_ONE
    idx = 0
    for d in ATE_LOOP_COUNT_NAF:
        f = fq12_sqr(f)
        for pc, p in pairs:
            f = _line(f, pc[1][idx], p)
        idx += 1
        if d != 0:
            for pc, p in pairs:
                f = _line(f, pc[1][idx], p)
            idx += 1
    for pc, p in pairs:
        f = _line(f, pc[1][idx], p)
    idx += 1
    for pc, p in pairs:
        f = _line(f, pc[1][idx], p)
    return f


def pairing(p, q):
    """groups::pairing, src/groups/mod.rs:894-902.  p, q Jacobian; infinity -> one."""
    pa = g_to_affine(FQ_OPS, p)
    qa = g_to_affine(FQ2_OPS, q)
    if pa is None or qa is None:
        return FQ12_ONE
    return fq12_final_exponentiation(miller_loop(g2_precompute(qa), pa))


def pairing_batch(ps, qs):
    """groups::pairing_batch, src/groups/mod.rs:904-926 (pairs with a point at infinity are skipped)."""
    pas, pcs = [], []
    for p, q in zip(ps, qs):
        pa = g_to_affine(FQ_OPS, p)
        qa = g_to_affine(FQ2_OPS, q)
        if pa is None or qa is None:
            continue
        pas.append(pa)
        pcs.append(g2_precompute(qa))
    if not pcs:
        return FQ12_ONE
    return fq12_final_exponentiation(miller_loop_batch(pcs, pas))


def lib_miller_loop_batch(pairs):
    """lib::miller_loop_batch, src/lib.rs:625-633: pairs are (G2, G1); infinity -> error (None here)."""
    pcs, pas = [], []
    for q, p in pairs:
        qa = g_to_affine(FQ2_OPS, q)
        if qa is None:
            return None
        pcs.append(g2_precompute(qa))
        pa = g_to_affine(FQ_OPS, p)
        if pa is None:
            return None
        pas.append(pa)
    return miller_loop_batch(pcs, pas)


# ---------------------------------------------------------------------------------------------
# Montgomery memory images (the crate's repr(C) layouts)
# ---------------------------------------------------------------------------------------------


def fq_to_mont(a):
    return (a * MONT_R) % Q


def fq_from_mont(m):
    return (m * pow(MONT_R, -1, Q)) % Q


def fr_to_mont(a):
    return (a * MONT_R) % R_ORDER


def fr_from_mont(m):
    return (m * pow(MONT_R, -1, R_ORDER)) % R_ORDER


def _flatten(x):
    if isinstance(x, int):
        return [x]
    out = []
    for c in x:
        out.extend(_flatten(c))
    return out


def to_bytes(x):
    """Any tower element / point (nested tuples of canonical Fq ints) -> little-endian Montgomery image."""
    return b"".join(fq_to_mont(c).to_bytes(32, "little") for c in _flatten(x))


def fq_list_from_bytes(b):
    assert len(b) % 32 == 0
    return [fq_from_mont(int.from_bytes(b[i:i + 32], "little")) for i in range(0, len(b), 32)]


def fq12_from_bytes(b):
    c = fq_list_from_bytes(b)
    assert len(c) == 12
    return (((c[0], c[1]), (c[2], c[3]), (c[4], c[5])), ((c[6], c[7]), (c[8], c[9]), (c[10], c[11])))


def g1_from_bytes(b):
    c = fq_list_from_bytes(b)
    return (c[0], c[1], c[2])


def g2_from_bytes(b):
    c = fq_list_from_bytes(b)
    return ((c[0], c[1]), (c[2], c[3]), (c[4], c[5]))


def fr_to_bytes(k):
    """repr(C) image of an Fr holding canonical scalar k (Montgomery mod r)."""
    return fr_to_mont(k).to_bytes(32, "little")


def gt_canonical_bytes(x):
    """Harness-defined canonical Gt serialisation: 12 x 32-byte big-endian canonical Fq."""
    return b"".join(c.to_bytes(32, "big") for c in _flatten(x))
