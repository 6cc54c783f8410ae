"""Python model of fq_inv_bgcd (csrc/bn_tower.cuh): binary extended Euclid in rounds of K = 30 steps on 64-bit approximations
(low 30 bits under the top 34 bits of the common bit window), the steps collected in a 2 x 2 matrix applied once per round to the
full-size values and, with an exact division by 2^K, to the cofactors modulo q.  Every bound the C code relies on is asserted
here (exact divisibility, |f| + |g| <= 2^K, the quotient range of the cofactor update, the end state (0, 1) after 17 rounds).
usage: bgcd_model.py [random cases]   (tests/test_hostsim.py runs the C code; tests/test_oracle_golden.py imports this model)"""
import random, sys
Q = 0x30644e72e131a029b85045b68181585d97816a916871ca8d3c208c16d87cfd47
K = 30
ROUNDS = 17
M64 = (1 << 64) - 1
QINV = pow(Q, -1, 1 << K)
NEGQINV = (-QINV) % (1 << K)

def approx(a, b):
    c = a | b
    limbs = [(c >> (32 * i)) & 0xffffffff for i in range(8)]
    j = max([i for i in range(8) if limbs[i]] + [0])
    if j <= 1:
        return a & M64, b & M64
    s = 32 - limbs[j].bit_length()           # clz of the top limb
    def top64(x):
        w = (x >> (32 * (j - 2))) & ((1 << 96) - 1)     # limbs j, j-1, j-2
        return ((w << s) >> 32) & M64
    ha, hb = top64(a), top64(b)
    lo = (1 << K) - 1
    return ((ha >> K) << K) | (a & lo), ((hb >> K) << K) | (b & lo)

def inner(a_, b_):
    f0, g0, f1, g1 = 1, 0, 0, 1
    mx = 0
    for _ in range(K):
        if a_ & 1:
            if a_ < b_:
                a_, b_ = b_, a_
                f0, f1 = f1, f0
                g0, g1 = g1, g0
            a_ = (a_ - b_) & M64
            f0 -= f1; g0 -= g1
        a_ >>= 1
        f1 <<= 1; g1 <<= 1
        mx = max(mx, abs(f0) + abs(g0), abs(f1) + abs(g1))
    assert mx <= 1 << K
    return f0, g0, f1, g1

def inv(y):
    a, b, u, v = y, Q, 1, 0
    for _ in range(ROUNDS):
        a_, b_ = approx(a, b)
        f0, g0, f1, g1 = inner(a_, b_)
        na = f0 * a + g0 * b
        nb = f1 * a + g1 * b
        assert na % (1 << K) == 0 and nb % (1 << K) == 0
        na >>= K; nb >>= K
        if na < 0: na, f0, g0 = -na, -f0, -g0
        if nb < 0: nb, f1, g1 = -nb, -f1, -g1
        assert na < (1 << 256) and nb < (1 << 256)
        def red(t):
            m = ((t & ((1 << K) - 1)) * NEGQINV) & ((1 << K) - 1)
            t = t + m * Q
            assert t % (1 << K) == 0
            t >>= K
            assert -Q < t < 2 * Q
            if t < 0: t += Q
            if t >= Q: t -= Q
            return t
        u, v = red(f0 * u + g0 * v), red(f1 * u + g1 * v)
        a, b = na, nb
    return a, b, v


def self_test(n_random=20000, seed=1):
    rng = random.Random(seed)
    vals = [1, 2, 3, Q - 1, Q - 2, (Q + 1) // 2, (Q - 1) // 2, 1 << 253, (1 << 253) + 1, (1 << 64) - 1, 1 << 64, (1 << 64) + 1, (1 << 128) - 1, 0xffffffff, 1 << 32]
    vals += [rng.randrange(1, Q) for _ in range(n_random)]
    vals += [rng.randrange(1, 1 << rng.randrange(1, 254)) for _ in range(n_random // 4)]
    vals += [Q - rng.randrange(1, 1 << rng.randrange(1, 200)) for _ in range(n_random // 4)]
    bad = 0
    for y in vals:
        a, b, v = inv(y)
        if not (a == 0 and b == 1 and v * y % Q == 1):
            bad += 1
    return len(vals), bad


if __name__ == "__main__":
    n, bad = self_test(int(sys.argv[1]) if len(sys.argv) > 1 else 20000)
    print("checked", n, "bad", bad)
    sys.exit(1 if bad else 0)
