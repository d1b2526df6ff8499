"""One process PER GPU (torchrun): contiguous shards and the single exchange step of the path.

Independent sponges never interact (reference: a sponge is a plain value mutated through `&mut self`,
src/poseidon/mod.rs:44-60), so batches shard with NO data-path collective: rank r owns the contiguous
slice shard_range(n, r, world) -- the library's own rule, psd_shard_range -- and runs the ordinary
single-GPU calls on it.  The only exchange is in the Merkle mode: every rank reduces its own subtree to
a root, the `world` roots (32 bytes each) are all-gathered over torch.distributed (NCCL over NVLink on
GPUs; gloo in the CPU tests) and the top log2(world) levels are hashed from them.

A single process that drives several GPUs does not need this module: psd_merkle_root_multi /
psd_permute_batch_multi (sponge_b200.MultiContext) do the same split inside the library.

Nothing here computes a hash itself.
"""
from __future__ import annotations

from typing import Optional

import numpy as np

from .poseidon import shard_range  # noqa: F401  (re-exported: the sharding rule lives in the library)


def _check_sizes(n_local: int, world: int) -> None:
    if n_local == 0 or n_local & (n_local - 1) or world & (world - 1):
        raise ValueError("leaves per rank and world size must be powers of two")


def distributed_merkle_root_dev(ctx, leaves_ptr: int, n_local: int, roots, top, stream: int = 0, group=None) -> None:
    """Device-resident form for one rank per GPU.  `ctx` is this rank's PoseidonContext; `leaves_ptr` its
    n_local leaves on its GPU; `roots` a (world + 1, 4) int64 CUDA tensor (row `world` receives this rank's
    subtree root, rows 0..world-1 the gathered ones), `top` a (1, 4) int64 CUDA tensor that receives the root
    of the whole tree on EVERY rank (the world - 1 top compressions are recomputed redundantly, which is
    cheaper than a broadcast).  Everything is queued on the current torch stream, which must be `stream`;
    nothing touches the host."""
    import torch.distributed as dist

    world = dist.get_world_size(group) if dist.is_initialized() else 1
    _check_sizes(n_local, world)
    if world == 1:
        ctx.merkle_root_dev(leaves_ptr, n_local, top.data_ptr(), stream=stream)
        return
    mine = roots[world:world + 1]
    ctx.merkle_root_dev(leaves_ptr, n_local, mine.data_ptr(), stream=stream)
    dist.all_gather_into_tensor(roots[:world], mine, group=group)   # world x 32 bytes: the only exchange
    ctx.merkle_root_dev(roots.data_ptr(), world, top.data_ptr(), stream=stream)


def distributed_merkle_root(compute, local_leaves: np.ndarray, group=None, device: Optional[str] = None) -> np.ndarray:
    """Host-array form: root of the tree whose leaves are the concatenation of every rank's `local_leaves`
    (rank order), returned on every rank.  `compute` is any object with
    `merkle_root(leaves: np.ndarray[(n,4) uint64]) -> np.ndarray[(4,) uint64]` -- in the product a
    sponge_b200.PoseidonContext."""
    import torch
    import torch.distributed as dist

    world = dist.get_world_size(group) if dist.is_initialized() else 1
    n_local = local_leaves.size // 4
    _check_sizes(n_local, world)
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
