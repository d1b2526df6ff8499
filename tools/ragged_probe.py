#!/usr/bin/env python3
"""psd_absorb_squeeze_ragged_batch through the host call, 2^20 BN254 t = 3 sponges, three length distributions, by
requested kernel variant (7: block-collective permutation with a vote; 5 / 3: warp-divergent integer kernels); results of
the three variants are compared with each other (the oracle comparison is tests/test_gpu_ragged.py)."""
import sys, os, time
import numpy as np
sys.path.insert(0, os.getcwd())
import sponge_b200 as S
from oracle import poseidon_ref as ref
ark, mds = ref.find_poseidon_ark_and_mds(ref.BN254_FR, 254, 2, 8, 57, 0)
cfg = S.PoseidonConfig(8, 57, 5, mds, ark, 2, 1, "bn254")
ctx = cfg.context()
rng = np.random.default_rng(1)
n = 1 << 20
for name, lens in (("uniform 0..18", rng.integers(0, 19, size=n)), ("all 8", np.full(n, 8)), ("1% long (200), rest 4", np.where(rng.random(n) < 0.01, 200, 4))):
    offsets = np.concatenate([[0], np.cumsum(lens)]).astype(np.uint64)
    x = np.zeros((int(offsets[-1]), 4), dtype=np.uint64); x[:, 0] = np.arange(len(x)) + 1
    out = {}
    res = {}
    for v in (7, 5, 3):
        ctx.set_variant(v)
        ts = []
        for it in range(4):
            t0 = time.perf_counter(); r = ctx.absorb_squeeze_ragged_batch(x, offsets, 1); ts.append(time.perf_counter() - t0)
        res[v] = r
        out[v] = round(min(ts) * 1e3, 1)
    assert np.array_equal(res[7], res[5]) and np.array_equal(res[7], res[3])
    print(name, "ms by requested variant:", out, flush=True)
