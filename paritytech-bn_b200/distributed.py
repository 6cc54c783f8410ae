"""One `pairing_batch` product sharded over several processes (one per GPU), SURVEY.md 8(e).

Independent pairings need no communication at all: rank d of D takes the contiguous index range
`shard_range(n, d, D)` and writes its own outputs.  The only exchange step exists when ONE product
prod_i e(P_i, Q_i) is split over ranks: every rank reduces its range to a single un-exponentiated
Fq12 (384 bytes, `Engine.miller_product`), the D partials are gathered (`torch.distributed.all_gather`
of a 384-byte tensor: NCCL over NVLink on GPUs, gloo in the CPU tests), multiplied on every rank in
rank order (<= D-1 Gt multiplications) and exponentiated ONCE.  Because Fq12 is a field and the
Miller value of a batch equals the product of the single Miller values exactly
(src/groups/mod.rs:609-640), the result is bit-identical to the single-device pairing_batch.

The in-process multi-GPU variant of the same protocol (peer copies instead of a process group) is
`pairing_product_sharded` in api.py / bn254_pairing_product_sharded in the C ABI.
"""
import numpy as np

from .api import G1_BYTES, G2_BYTES, GT_BYTES


def shard_range(n, rank, world):
    """Contiguous index range [lo, hi) of rank `rank` out of `world` over n units."""
    if world <= 0 or not (0 <= rank < world):
        raise ValueError("bad rank/world")
    return n * rank // world, n * (rank + 1) // world


def combine_partials(engine, partials):
    """Product of the gathered 384-byte Miller partials (rank order) and ONE final exponentiation.
    Returns (Gt bytes, is_one)."""
    partials = np.ascontiguousarray(partials, dtype=np.uint8).reshape(-1, GT_BYTES)
    acc = partials[0].copy()
    for d in range(1, partials.shape[0]):
        acc = engine.gt_mul(acc, partials[d])
    out, some = engine.final_exponentiation(acc)
    if not some[0]:
        raise ArithmeticError("miller loop cannot produce zero")   # the reference's expect(), src/groups/mod.rs:925
    one = engine.gt_one()
    return out, out.tobytes() == one.tobytes()


def pairing_batch_distributed(engine, g1, g2, group=None, device=None):
    """pairing_batch over the union of all ranks' pairs.  `g1`/`g2` hold THIS rank's shard (any length,
    including empty).  Every rank returns the same (Gt bytes, is_one)."""
    import torch
    import torch.distributed as dist
    g1 = np.ascontiguousarray(g1, dtype=np.uint8).reshape(-1)
    g2 = np.ascontiguousarray(g2, dtype=np.uint8).reshape(-1)
    assert g1.size % G1_BYTES == 0 and g2.size % G2_BYTES == 0 and g1.size // G1_BYTES == g2.size // G2_BYTES
    partial = engine.miller_product(g1, g2)                      # 384 bytes, this rank's un-exponentiated product
    world = dist.get_world_size(group)
    t = torch.from_numpy(partial.copy())
    if device is not None:
        t = t.to(device)
    gathered = [torch.empty_like(t) for _ in range(world)]
    dist.all_gather(gathered, t, group=group)                    # D x 384 bytes: latency-bound, no per-pairing traffic
    partials = np.stack([x.cpu().numpy() for x in gathered])
    return combine_partials(engine, partials)
