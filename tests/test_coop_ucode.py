"""The microcode of the warp-cooperative pairing (tools/gen_coop_ucode.py -> csrc/bn_coop_ucode.cuh) on the CPU: the
generator's simulator runs the very tables the GPU executes, on Python integers, against the oracle."""
import os
import random
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tools"))
import gen_coop_ucode as g  # noqa: E402

from oracle import bn254_ref as o  # noqa: E402

rng = random.Random(77)


def _pair():
    p = o.g_mul(o.FQ_OPS, o.G1_ONE, rng.randrange(1, o.R_ORDER))
    q = o.g_mul(o.FQ2_OPS, o.G2_ONE, rng.randrange(1, o.R_ORDER))
    return p, q


def test_miller_and_pairing_match_the_oracle():
    for _ in range(2):
        p, q = _pair()
        pa, qa = o.g_to_affine(o.FQ_OPS, p), o.g_to_affine(o.FQ2_OPS, q)
        assert g.simulate(p, q, "miller") == o.miller_loop(o.g2_precompute(qa), pa)
        assert g.simulate(p, q) == o.pairing(p, q)
    # z == 1 inputs (the generators) go the same way
    assert g.simulate(o.G1_ONE, o.G2_ONE) == o.pairing(o.G1_ONE, o.G2_ONE)


def test_paired_program_of_the_quad_lane_executor():
    """The line walk on lanes 24..31 beside the squarings / line multiplications on lanes 0..23 (one stage ahead, lines through
    four buffers): same Miller value and pairing as the oracle, from the routines (with the cross-half hazard check: neither
    half of a paired step reads or writes what the other writes) and from the ENCODED two-word trace."""
    p, q = _pair()
    pa, qa = o.g_to_affine(o.FQ_OPS, p), o.g_to_affine(o.FQ2_OPS, q)
    assert g.simulate_paired(p, q, "miller") == o.miller_loop(o.g2_precompute(qa), pa)
    assert g.simulate_paired(p, q) == o.pairing(p, q) == g.decode_and_run_paired(p, q)
    assert g.simulate_paired(o.G1_ONE, o.G2_ONE) == o.pairing(o.G1_ONE, o.G2_ONE)
    steps, trace, marks, nseq = g.flatten4()
    seq_steps, seq_trace, _ = g.flatten()
    assert len(trace) < 0.66 * len(seq_trace)                     # 1,072 entries against 1,651 steps
    assert all(s < nseq for s, _, _, _ in seq_trace)              # the one-warp executor loads a prefix of the step table
    assert all(len(steps[sw].lanes) <= 8 and (sf is None or len(steps[sf].lanes) <= 24) for sf, sw, _, _, _ in trace if sw is not None)


def test_final_exponentiation_phase_alone():
    p, q = _pair()
    f = g.simulate(p, q, "miller")
    assert g.simulate_fexp(f) == o.fq12_final_exponentiation(f)


def test_pow_walk_with_in_place_routines():
    """SQR12 on the accumulator and MUL12 with D aliasing A, as k_coop_gt_pow calls them (Gt::pow, src/lib.rs:592)."""
    p, q = _pair()
    a = g.simulate(p, q)
    for k in (0, 1, 2, 0xb5, rng.randrange(1 << 70)):
        assert g.simulate_pow(a, k) == o.fq12_pow(a, k)
        assert g.simulate_pow4(a, k) == o.fq12_pow(a, k)           # the quad-lane executor's SQR12B / MUL12T / T0INIT


def test_encoded_tables_decode_to_the_same_result():
    """What is emitted (16-bit terms, lane words, headers, trace) is what was simulated."""
    p, q = _pair()
    assert g.decode_and_run(p, q) == o.pairing(p, q)


def test_committed_header_is_up_to_date(tmp_path):
    out = tmp_path / "ucode.cuh"
    g.emit(str(out))
    assert out.read_text() == open(os.path.join(ROOT, "paritytech-bn_b200", "csrc", "bn_coop_ucode.cuh")).read()


def test_window_digits_reconstruct_u():
    d = g.wnaf4(o.BN_U)
    assert sum(x << i for i, x in enumerate(d)) == o.BN_U and all(x == 0 or (x % 2 and -7 <= x <= 7) for x in d)


def test_tables_and_slot_files_fit_the_shared_memory_of_one_sm():
    """The launch shapes of csrc/bn_coop.cuh (16 warps per block) and csrc/bn_coop4.cuh (4 groups per block) against the 227 KB a
    block may use on sm_100a, from the generated sizes: a routine added to the generator must not silently break a launch."""
    import re
    hdr = open(os.path.join(ROOT, "paritytech-bn_b200", "csrc", "bn_coop_ucode.cuh")).read()
    d = {k: int(v) for k, v in re.findall(r"#define (BN_COOP4?_[A-Z_0-9]+) (\d+)", hdr)}
    slot_bytes = d["BN_COOP_NSLOTS"] * 64
    seq_words = (d["BN_COOP_NSTEPS_SEQ"] * 160 + d["BN_COOP_NSTEPS_SEQ"] + d["BN_COOP_NTRACE"] + 3) & ~3
    quad_words = (d["BN_COOP_NSTEPS"] * 160 + d["BN_COOP_NSTEPS"] + 2 * d["BN_COOP4_NTRACE"] + 3) & ~3
    limit = 227 * 1024
    assert seq_words * 4 + 16 * slot_bytes <= limit, (seq_words * 4, slot_bytes)
    assert quad_words * 4 + 4 * slot_bytes <= limit
    src = open(os.path.join(ROOT, "paritytech-bn_b200", "csrc", "bn_coop.cuh")).read()
    src4 = open(os.path.join(ROOT, "paritytech-bn_b200", "csrc", "bn_coop4.cuh")).read()
    assert "#define BN_COOP_MAX_WARPS 16" in src and "#define BN_COOP4_MAX_GROUPS 4" in src4
    assert d["BN_COOP4_W_LANE0"] == 24 and d["BN_COOP_NSTEPS"] < 255          # 0xff is "no step" in the paired trace
