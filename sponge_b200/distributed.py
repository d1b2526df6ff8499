"""Multi-GPU driver: one process per GPU, contiguous shards, and the single exchange step of the path.

Independent sponges never interact (reference: a sponge is a plain value mutated through `&mut self`,
src/poseidon/mod.rs:44-60), so batches shard with NO data-path collective: rank r owns the contiguous
slice shard_range(n, r, world) and runs the ordinary single-GPU calls on it.  The only exchange is in
the Merkle mode: every rank reduces its own subtree to a root, the `world` roots (32 bytes each) are
all-gathered over torch.distributed (NCCL over NVLink on GPUs; gloo in the CPU tests) and the top
log2(world) levels are hashed from them.

`compute` is any object with `merkle_root(leaves: np.ndarray[(n,4) uint64]) -> np.ndarray[(4,) uint64]`
-- in the product a sponge_b200.PoseidonContext (GPU).  Nothing here computes a hash itself.
"""
from __future__ import annotations

from typing import Optional, Tuple

import numpy as np


def shard_range(n_total: int, rank: int, world: int) -> Tuple[int, int]:
    """Contiguous shard [lo, hi) of rank `rank`; the first n_total % world ranks get one extra unit."""
    if not 0 <= rank < world:
        raise ValueError("rank outside [0, world)")
    base, extra = divmod(n_total, world)
    lo = rank * base + min(rank, extra)
    return lo, lo + base + (1 if rank < extra else 0)


def distributed_merkle_root(compute, local_leaves: np.ndarray, group=None, device: Optional[str] = None) -> np.ndarray:
    """Root of the tree whose leaves are the concatenation of every rank's `local_leaves` (rank order).

    Each rank must hold the same power-of-two number of leaves and `world` must be a power of two.
    Returns the root on every rank (the top levels are recomputed redundantly from the gathered
    roots -- `world - 1` compressions -- which is cheaper than a broadcast)."""
    import torch
    import torch.distributed as dist

    world = dist.get_world_size(group) if dist.is_initialized() else 1
    n_local = local_leaves.size // 4
    if n_local == 0 or n_local & (n_local - 1) or world & (world - 1):
        raise ValueError("leaves per rank and world size must be powers of two")
    local_root = np.ascontiguousarray(compute.merkle_root(local_leaves), dtype=np.uint64).reshape(4)
    if world == 1:
        return local_root
    if device is None:
        device = "cuda" if dist.get_backend(group) == "nccl" else "cpu"
    mine = torch.from_numpy(local_root.view(np.int64).copy()).to(device)
    gathered = torch.empty((world, 4), dtype=torch.int64, device=device)
    dist.all_gather_into_tensor(gathered, mine.reshape(1, 4), group=group)   # world x 32 bytes: the only exchange
    roots = gathered.cpu().numpy().view(np.uint64)
    return np.ascontiguousarray(compute.merkle_root(np.ascontiguousarray(roots)), dtype=np.uint64).reshape(4)
