"""world_size-2 `gloo` test of the N>1 host logic (no GPU): index-range sharding, the 384-byte partial
gather and the combine step of `pairing_batch_distributed`.  The CUDA engine is replaced by an
ORACLE-BACKED STUB here (tests may use the oracle as the checker); the real engine runs the same
protocol in tests/test_gpu_parity.py::test_distributed_protocol_single_process."""
import os
import subprocess
import sys
import textwrap

import numpy as np

import paritytech_bn_b200 as bn

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_shard_ranges_cover_and_are_contiguous():
    for n in (0, 1, 7, 64, 65536, (1 << 22) + 3):
        for world in (1, 2, 3, 4, 8):
            r = [bn.shard_range(n, d, world) for d in range(world)]
            assert r[0][0] == 0 and r[-1][1] == n
            assert all(r[d][1] == r[d + 1][0] for d in range(world - 1))
            sizes = [hi - lo for lo, hi in r]
            assert max(sizes) - min(sizes) <= 1


def test_cancelling_pairs_construction_shards_for_any_world_size():
    """bench.py's sharded_product (BASELINE configs[4]): the per-rank slices of the known-answer construction tile the whole
    pair list for even AND odd world sizes (ranges crossing the half-way point), the scalars cancel (sum a_i b_i = 0 mod r),
    and -- on a tiny instance through the oracle -- the product of all pairs is one while a rank's partial is not."""
    import paritytech_bn_b200 as bn
    from oracle import c_oracle as c
    from tests import util
    half = 9
    A, B = util.fr_images(half, stream=401), util.fr_images(half, stream=402)
    n = 2 * half
    full1, full2 = util.cancelling_pair_scalars(A, B, 0, n)
    rinv = pow(1 << 256, -1, util.R)

    def canon(img):
        return [int.from_bytes(img[i:i + 32].tobytes(), "little") * rinv % util.R for i in range(0, len(img), 32)]
    a, b = canon(full1), canon(full2)
    assert len(a) == n and sum(x * y for x, y in zip(a, b)) % util.R == 0
    for world in (1, 2, 3, 4, 5, 7, 8):
        parts = [util.cancelling_pair_scalars(A, B, *bn.shard_range(n, r, world)) for r in range(world)]
        assert np.concatenate([p[0] for p in parts]).tobytes() == full1.tobytes()
        assert np.concatenate([p[1] for p in parts]).tobytes() == full2.tobytes()
    e1, e2 = util.cancelling_pair_scalars(A, B, 5, 5)
    assert e1.size == 0 and e2.size == 0
    g1 = c.g1_mul(np.tile(util.G1_GEN, n), full1)
    g2 = c.g2_mul(np.tile(util.G2_GEN, n), full2)
    assert c.pairing_batch(g1, g2).tobytes() == util.GT_ONE.tobytes()
    lo, hi = bn.shard_range(n, 1, 3)
    assert c.pairing_batch(g1[lo * 96:hi * 96], g2[lo * 192:hi * 192]).tobytes() != util.GT_ONE.tobytes()


WORKER = textwrap.dedent('''
    import os, sys
    sys.path.insert(0, %(root)r)
    import numpy as np
    import torch.distributed as dist
    import paritytech_bn_b200 as bn
    from oracle import c_oracle as c
    from tests import util

    class OracleEngine:   # test-only stand-in for bn.Engine: same method names, CPU oracle underneath
        def miller_product(self, g1, g2):
            return c.miller_loop_batch_skip(g2, g1)
        def gt_mul(self, a, b):
            return c.fq12_mul(a, b)
        def final_exponentiation(self, a):
            return c.final_exp(a)
        def gt_one(self):
            return util.GT_ONE

    dist.init_process_group("gloo")
    rank, world = dist.get_rank(), dist.get_world_size()
    n = 9                                     # odd: ranks get ragged shards (4 and 5)
    ks1, ks2 = util.scalars(n, stream=51), util.scalars(n, stream=52)
    g1, g2 = util.points_cpu(ks1, ks2)
    g1 = g1.reshape(n, 96).copy(); g1[3] = util.G1_INF; g1 = g1.reshape(-1)     # a skipped pair
    lo, hi = bn.shard_range(n, rank, world)
    out, is_one = bn.pairing_batch_distributed(OracleEngine(), g1[lo * 96:hi * 96], g2[lo * 192:hi * 192])
    exp = c.pairing_batch(g1, g2)
    assert out.tobytes() == exp.tobytes(), "sharded product differs from the single-process pairing_batch"
    assert not is_one
    # an empty shard on rank 1 and everything on rank 0
    mine = slice(0, n) if rank == 0 else slice(0, 0)
    out2, _ = bn.pairing_batch_distributed(OracleEngine(), g1.reshape(n, 96)[mine].reshape(-1), g2.reshape(n, 192)[mine].reshape(-1))
    assert out2.tobytes() == exp.tobytes()
    dist.destroy_process_group()
    print("rank", rank, "ok")
''')


def _free_port():
    import socket
    with socket.socket() as sk:
        sk.bind(("127.0.0.1", 0))
        return str(sk.getsockname()[1])


def test_pairing_batch_distributed_gloo_world2(tmp_path):
    script = tmp_path / "worker.py"
    script.write_text(WORKER % {"root": ROOT})
    env = dict(os.environ, MASTER_ADDR="127.0.0.1", OMP_NUM_THREADS="2")
    p = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
                        "--master-port", _free_port(), str(script)], capture_output=True, text=True, timeout=600, env=env, cwd=ROOT)
    assert p.returncode == 0, p.stdout[-2000:] + p.stderr[-4000:]
    assert "rank 0 ok" in p.stdout and "rank 1 ok" in p.stdout


def test_reference_arm_under_torchrun_prints_one_line(tmp_path):
    """bench.py --impl reference at N=2: rank 0 alone runs and prints, the other rank exits 0 silently."""
    env = dict(os.environ, MASTER_ADDR="127.0.0.1")
    p = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
                        "--master-port", _free_port(), os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "2", "--steps", "1",
                        "--warmup", "0", "--batch-log2", "6"], capture_output=True, text=True, timeout=900, env=env, cwd=ROOT)
    assert p.returncode == 0, p.stderr[-3000:]
    lines = [ln for ln in p.stdout.splitlines() if ln.startswith("{")]
    assert len(lines) == 1
    import json
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["cpu_baseline"]["kind"] == "port" and d["e2e"]["h2d_bytes_per_step"] == 0
