#!/usr/bin/env python3
"""Time of ONE launch of n BN254 t = 3 permutations (device-resident, CUDA events, median of 8) for n = 32 .. 262 144 and
kernel variants 7, 6, 5, 3: below one wave the time is the latency of a single permutation -- what every small level of a
Merkle tree costs."""
import json, sys, os
import numpy as np
sys.path.insert(0, os.getcwd())
import torch
import sponge_b200 as S
from oracle import poseidon_ref as ref
ark, mds = ref.find_poseidon_ark_and_mds(ref.BN254_FR, 254, 2, 8, 57, 0)
cfg = S.PoseidonConfig(8, 57, 5, mds, ark, 2, 1, "bn254")
ctx = cfg.context()
x = torch.empty((1 << 20) * 3, 4, dtype=torch.int64, device="cuda:0")
ctx.synth_fill_dev(x.data_ptr(), (1 << 20) * 3, 1)
for n in (32, 128, 1024, 8192, 18944, 65536, 262144):
    out = {}
    for v in (7, 6, 5, 3):
        ctx.set_variant(v)
        ms = []
        for it in range(12):
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record(); ctx.permute_batch_dev(x.data_ptr(), n); b.record(); torch.cuda.synchronize()
            if it >= 4: ms.append(a.elapsed_time(b))
        out[f"v{ctx.active_variant()}"] = round(float(np.median(ms)) * 1e3, 1)
    print(n, out, flush=True)
