"""The N>1 path on CPU: world_size 2 over gloo.  The distributed driver (sponge_b200/distributed.py)
only shards and gathers; here its `compute` is an oracle-backed stand-in (tests only), so the test
checks the sharding arithmetic, the synthetic-index offsets bench.py uses per rank, and the root
gather + top levels of the Merkle mode against a single-process oracle over all leaves."""
import os
import socket

import numpy as np
import pytest

from conftest import oracle_cfg


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


class OracleCompute:
    """Stand-in for PoseidonContext in CPU tests (the product never uses the oracle)."""

    def __init__(self):
        from oracle import poseidon_ref as R
        from oracle.c_oracle import COracle
        self.orc = COracle(oracle_cfg(R, "bn254_t3"), 0)

    def merkle_root(self, leaves):
        return self.orc.merkle_root(np.ascontiguousarray(leaves))


def _worker(rank, world, port, n_local, out_dir):
    import sys
    sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__))))
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    import torch.distributed as dist
    from sponge_b200.distributed import distributed_merkle_root, shard_range
    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
    try:
        comp = OracleCompute()
        lo, hi = shard_range(n_local * world, rank, world)
        assert hi - lo == n_local
        leaves = comp.orc.synth(n_local, seed=0xC4, first=lo)        # same index space as bench.py: first = rank * n
        root = distributed_merkle_root(comp, leaves)
        # independent-batch mode: each rank permutes its own shard, nothing is exchanged
        states = comp.orc.synth(3 * 64, seed=0xC2, first=3 * 64 * rank)
        np.save(os.path.join(out_dir, f"root_{rank}.npy"), root)
        np.save(os.path.join(out_dir, f"perm_{rank}.npy"), comp.orc.permute_batch(states))
        dist.barrier()
    finally:
        dist.destroy_process_group()


def test_shard_range_partitions_exactly():
    from sponge_b200.distributed import shard_range
    for n in (0, 1, 7, 8, 1 << 24, (1 << 24) + 5):
        for world in (1, 2, 3, 4, 8):
            spans = [shard_range(n, r, world) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == n
            assert all(spans[i][1] == spans[i + 1][0] for i in range(world - 1))
            sizes = [b - a for a, b in spans]
            assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        shard_range(8, 2, 2)


def test_two_rank_merkle_and_sharded_permute(tmp_path):
    import torch.multiprocessing as mp
    world, n_local = 2, 256
    port = _free_port()
    mp.spawn(_worker, args=(world, port, n_local, str(tmp_path)), nprocs=world, join=True)
    comp = OracleCompute()
    all_leaves = comp.orc.synth(n_local * world, seed=0xC4, first=0)
    want = comp.orc.merkle_root(all_leaves)
    for r in range(world):
        assert np.array_equal(np.load(tmp_path / f"root_{r}.npy"), want)
    all_states = comp.orc.synth(3 * 64 * world, seed=0xC2, first=0)
    want_perm = comp.orc.permute_batch(all_states.copy())
    got = np.concatenate([np.load(tmp_path / f"perm_{r}.npy") for r in range(world)])
    assert np.array_equal(got, want_perm)
