"""Seeded workload construction shared by the tests and bench.py.

Scalars come from SplitMix64 (seed 0x424E323534 + stream id), four 64-bit words reduced mod r, zero
rejected.  Points are scalar multiples of the generators, kept as Jacobian triples with z != 1 (what
`G * Fr` returns and what the reference's `pairing` normally receives).  CPU-side point generation
uses the C oracle (test infrastructure); bench.py generates its large inputs with the CUDA
scalar-multiplication kernels instead.
"""
import numpy as np

Q = 0x30644e72e131a029b85045b68181585d97816a916871ca8d3c208c16d87cfd47
R = 0x30644e72e131a029b85045b68181585d2833e84879b9709143e1f593f0000001
MONT = 1 << 256
SEED = 0x424E323534
_M64 = (1 << 64) - 1


def splitmix64_stream(seed, n):
    """n 64-bit outputs of SplitMix64 as a numpy uint64 array (vectorised)."""
    with np.errstate(over="ignore"):
        idx = np.arange(1, n + 1, dtype=np.uint64)
        z = np.uint64(seed & _M64) + idx * np.uint64(0x9E3779B97F4A7C15)
        z = (z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
        z = (z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
        return z ^ (z >> np.uint64(31))


def scalars(n, stream=0):
    """n canonical scalars in [1, r) as Python ints."""
    w = splitmix64_stream(SEED + 0x1000003 * stream, 4 * n).reshape(n, 4)
    out = []
    for row in w:
        k = (int(row[0]) | (int(row[1]) << 64) | (int(row[2]) << 128) | (int(row[3]) << 192)) % R
        out.append(k if k else 1)
    return out


def fr_bytes(ks):
    """repr(C) Fr images (Montgomery mod r) of canonical scalars."""
    return np.frombuffer(b"".join(((k % R) * MONT % R).to_bytes(32, "little") for k in ks), dtype=np.uint8).copy()


def fr_images(n, stream=0):
    """n random repr(C) Fr images without Python big integers (bench-size inputs): four SplitMix64 words each with the top
    word cut to 61 bits, so every image is below 2^253 < r.  A uniform image is a uniform scalar (Montgomery scaling is a
    bijection of Z_r); zero is avoided by setting the lowest bit."""
    w = splitmix64_stream(SEED + 0x1000003 * stream + 0x5bd1e995, 4 * n).reshape(n, 4).copy()
    w[:, 3] &= np.uint64((1 << 61) - 1)
    w[:, 0] |= np.uint64(1)
    return w.view(np.uint8).reshape(-1).copy()


def neg_fr_images(fr):
    """Images of the negated scalars: r - x limb-wise (numpy multiword borrow arithmetic); x must be in (0, r)."""
    x = np.ascontiguousarray(fr, dtype=np.uint8).view(np.uint64).reshape(-1, 4)
    out = np.empty_like(x)
    borrow = np.zeros(x.shape[0], dtype=np.uint64)
    with np.errstate(over="ignore"):
        for j in range(4):
            rj = np.uint64((R >> (64 * j)) & _M64)
            t = rj - x[:, j]
            b1 = (x[:, j] > rj).astype(np.uint64)
            t2 = t - borrow
            b2 = (borrow > t).astype(np.uint64)
            out[:, j] = t2
            borrow = b1 | b2
    assert not borrow.any()
    return out.view(np.uint8).reshape(-1).copy()


def cancelling_pair_scalars(A, B, i0, i1):
    """Scalar images of the global pair indices [i0, i1) of the known-answer product of SURVEY 8e, n = 2 * len(A) pairs:
    pair i < n/2 is (A_i G1, B_i G2), pair n/2 + j is ((-B_j) G1, A_j G2), so sum a_i b_i = 0 mod r and the whole product is
    exactly one while any proper sub-range is not.  Returns (G1 scalar images, G2 scalar images) of the range."""
    half = len(A) // 32
    assert len(B) == len(A) and 0 <= i0 <= i1 <= 2 * half
    s1, s2 = [np.zeros(0, dtype=np.uint8)], [np.zeros(0, dtype=np.uint8)]
    if i0 < half:
        e = min(i1, half)
        s1.append(A[i0 * 32:e * 32]); s2.append(B[i0 * 32:e * 32])
    if i1 > half:
        b = max(i0, half) - half
        s1.append(neg_fr_images(B[b * 32:(i1 - half) * 32])); s2.append(A[b * 32:(i1 - half) * 32])
    return np.concatenate(s1), np.concatenate(s2)


def fq_mont_bytes(vals):
    return b"".join((v * MONT % Q).to_bytes(32, "little") for v in vals)


G1_GEN = np.frombuffer(fq_mont_bytes([1, 2, 1]), dtype=np.uint8).copy()
G2_GEN = np.frombuffer(fq_mont_bytes([
    10857046999023057135944570762232829481370756359578518086990519993285655852781,
    11559732032986387107991004021392285783925812861821192530917403151452391805634,
    8495653923123431417604973247489272438418190587263600148770280649306958101930,
    4082367875863433681332203403145435568316851327593401208105741076214120093531, 1, 0]), dtype=np.uint8).copy()
G1_INF = np.frombuffer(fq_mont_bytes([0, 1, 0]), dtype=np.uint8).copy()
G2_INF = np.frombuffer(fq_mont_bytes([0, 0, 1, 0, 0, 0]), dtype=np.uint8).copy()
GT_ONE = np.frombuffer(fq_mont_bytes([1] + [0] * 11), dtype=np.uint8).copy()


def neg_g1(p):
    """-P on G1 Montgomery images (y -> q - y; infinity untouched), src/groups/mod.rs:336-350."""
    p = np.array(p, dtype=np.uint8).reshape(-1, 96).copy()
    for row in p:
        if not row[64:96].any():
            continue
        y = int.from_bytes(row[32:64].tobytes(), "little")
        row[32:64] = np.frombuffer(((Q - y) % Q).to_bytes(32, "little"), dtype=np.uint8)
    return p.reshape(-1)


def points_cpu(ks1, ks2):
    """(a_i * G1, b_i * G2) as Jacobian images, via the C oracle."""
    from oracle import c_oracle as c
    g1 = c.g1_mul(np.tile(G1_GEN, len(ks1)), fr_bytes(ks1))
    g2 = c.g2_mul(np.tile(G2_GEN, len(ks2)), fr_bytes(ks2))
    return g1, g2


def groth16_checks(n_checks, stream=7, scalar_mul=None):
    """Config #3: synthetic Groth16-shaped 4-pair checks with one shared verifying key.

    Pairs per check: (a*G1, b*G2), (-alpha*G1, beta*G2), (-l*G1, gamma*G2), (-c*G1, delta*G2) with
    c = (a*b - alpha*beta - l*gamma)/delta mod r, so the product is one.  Every second check is made
    invalid by adding 1 to c.  Returns dict with the flat 4-pair layout, the split layout for the
    shared-table entry point, and the expected accept bytes.
    scalar_mul(gen_name, ks) -> Jacobian images lets bench.py use the CUDA kernels.
    """
    if scalar_mul is None:
        def scalar_mul(which, ks):
            from oracle import c_oracle as c
            if which == "g1":
                return c.g1_mul(np.tile(G1_GEN, len(ks)), fr_bytes(ks))
            return c.g2_mul(np.tile(G2_GEN, len(ks)), fr_bytes(ks))
    alpha, beta, gamma, delta = scalars(4, stream=stream)
    a = scalars(n_checks, stream=stream + 1)
    b = scalars(n_checks, stream=stream + 2)
    ll = scalars(n_checks, stream=stream + 3)
    dinv = pow(delta, -1, R)
    cs, valid = [], []
    for i in range(n_checks):
        c = (a[i] * b[i] - alpha * beta - ll[i] * gamma) * dinv % R
        ok = (i % 2 == 0)
        if not ok:
            c = (c + 1) % R
        cs.append(c)
        valid.append(ok)
    g1_a = scalar_mul("g1", a).reshape(n_checks, 96)
    g2_b = scalar_mul("g2", b).reshape(n_checks, 192)
    g1_alpha = neg_g1(scalar_mul("g1", [alpha]))
    g1_l = neg_g1(scalar_mul("g1", ll)).reshape(n_checks, 96)
    g1_c = neg_g1(scalar_mul("g1", [c if c else 1 for c in cs])).reshape(n_checks, 96)
    vk = scalar_mul("g2", [beta, gamma, delta]).reshape(3, 192)
    g1_flat = np.empty((n_checks, 4, 96), dtype=np.uint8)
    g2_flat = np.empty((n_checks, 4, 192), dtype=np.uint8)
    g1_flat[:, 0], g1_flat[:, 1], g1_flat[:, 2], g1_flat[:, 3] = g1_a, g1_alpha, g1_l, g1_c
    g2_flat[:, 0], g2_flat[:, 1], g2_flat[:, 2], g2_flat[:, 3] = g2_b, vk[0], vk[1], vk[2]
    return {
        "g1": g1_flat.reshape(-1), "g2": g2_flat.reshape(-1),
        "vk_g2": vk.reshape(-1),
        "g1_shared": np.ascontiguousarray(g1_flat[:, 1:4]).reshape(-1),
        "g1_var": np.ascontiguousarray(g1_flat[:, 0]).reshape(-1),
        "g2_var": np.ascontiguousarray(g2_flat[:, 0]).reshape(-1),
        "valid": np.array(valid, dtype=np.uint8),
    }


# ---------------------------------------------------------------------------------------------
# SURVEY 8f N2/N3 inputs: affine points for AffineG::new and compressed encodings for from_compressed
# ---------------------------------------------------------------------------------------------
def affine_of(jac_images, unit):
    """(x, y) Montgomery images of normalised Jacobian images (z = one); unit = 96 (G1) or 192 (G2)."""
    a = np.asarray(jac_images, dtype=np.uint8).reshape(-1, unit)
    return a[:, : 2 * unit // 3].reshape(-1).copy()


def _rand_fq(rng):
    return int.from_bytes(rng.bytes(40), "little") % Q


def g2_curve_points_outside_subgroup(n, seed=11):
    """n affine points of the twist E'(Fq2) found by solving y^2 = x^3 + b' for random x.  The group has order
    r * (2q - r), so such a point lies in the r-torsion only with probability ~2^-254: AffineG2::new must
    answer NotInSubgroup.  Canonical ((x0, x1), (y0, y1)) tuples."""
    from oracle import bn254_ref as o
    rng = np.random.default_rng(seed)
    out = []
    while len(out) < n:
        x = (_rand_fq(rng), _rand_fq(rng))
        y = o.fq2_sqrt(o.fq2_add(o.fq2_mul(o.fq2_sqr(x), x), o.G2_B))
        if y is not None:
            out.append((x, y))
    return out


def g2_small_order_points(n, seed=12):
    """Points of E'(Fq2) with no r-torsion component: r * P for curve points P (order divides the cofactor
    2q - r).  The hardest negatives for any shortcut subgroup test."""
    from oracle import bn254_ref as o
    out = []
    for x, y in g2_curve_points_outside_subgroup(n, seed):
        t = o.g_to_affine(o.FQ2_OPS, o.g_mul(o.FQ2_OPS, (x, y, o.FQ2_ONE), o.R_ORDER))
        if t is not None:
            out.append(t)
    return out


def affine_bytes(points):
    """Montgomery images of canonical affine tuples: ints for G1 coordinates, pairs for G2 coordinates."""
    flat = []
    for x, y in points:
        for c in (x, y):
            flat.extend(c if isinstance(c, tuple) else (c,))
    return np.frombuffer(fq_mont_bytes(flat), dtype=np.uint8).copy()


def compress_g1(x, y):
    """Inverse of G1::from_compressed (src/lib.rs:359-375): tag 2 for even canonical y, 3 for odd, x big-endian."""
    return bytes([3 if y & 1 else 2]) + x.to_bytes(32, "big")


def compress_g2(x, y):
    """Inverse of G2::from_compressed (src/lib.rs:506-526): tag 10 when y is the smaller of (y, -y) as the
    integer c1 * q + c0, else 11; x as the 512-bit big-endian integer c1 * q + c0."""
    yn = ((-y[0]) % Q, (-y[1]) % Q)
    gt = y[1] * Q + y[0] > yn[1] * Q + yn[0]
    return bytes([11 if gt else 10]) + (x[1] * Q + x[0]).to_bytes(64, "big")
