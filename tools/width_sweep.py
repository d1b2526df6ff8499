#!/usr/bin/env python3
"""Permutations per second by width and kernel variant on one GPU (device-resident states, CUDA events, 2^22 sponges,
median of 5 launches after 2 warm-ups).  Parameter sets: the oracle's generator (oracle/poseidon_ref.py), the round
counts of the reference's default tables.  Output: one JSON line per (field, t, alpha)."""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import torch
    import sponge_b200 as S
    from oracle import poseidon_ref as ref
    n = 1 << 22
    shapes = [("bn254", ref.BN254_FR, 254, 1, 8, 56, 5), ("bn254", ref.BN254_FR, 254, 2, 8, 57, 5),
              ("bn254", ref.BN254_FR, 254, 3, 8, 56, 5), ("bn254", ref.BN254_FR, 254, 4, 8, 60, 5),
              ("bls12_381", ref.BLS12_381_FR, 255, 2, 8, 29, 17), ("bls12_381", ref.BLS12_381_FR, 255, 2, 8, 56, 5),
              ("bls12_381", ref.BLS12_381_FR, 255, 3, 8, 56, 5), ("bls12_381", ref.BLS12_381_FR, 255, 4, 8, 56, 5)]
    for fname, p, bits, rate, rf, rp, alpha in shapes:
        ark, mds = ref.find_poseidon_ark_and_mds(p, bits, rate, rf, rp, 0)
        t = rate + 1
        cfg = S.PoseidonConfig(rf, rp, alpha, mds, ark, rate, 1, fname)
        ctx = cfg.context()
        x = torch.empty((n * t, 4), dtype=torch.int64, device="cuda:0")
        ctx.synth_fill_dev(x.data_ptr(), n * t, 0x5EED)
        rec = dict(field=fname, t=t, alpha=alpha, rf=rf, rp=rp, n=n, perms_per_s={})
        for variant in (7, 6, 3):
            ctx.set_variant(variant)
            active = ctx.active_variant()
            ms = []
            for it in range(7):
                a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                a.record()
                ctx.permute_batch_dev(x.data_ptr(), n)
                b.record()
                torch.cuda.synchronize()
                if it >= 2:
                    ms.append(a.elapsed_time(b))
            rec["perms_per_s"][f"requested_{variant}"] = dict(active=active, value=n / (float(np.median(ms)) * 1e-3))
        ctx.close()
        print(json.dumps(rec), flush=True)


if __name__ == "__main__":
    main()
