"""Variant-7 kernels of DIFFERENT shapes running at the same time on one GPU.  Every block of that kernel allocates tensor
memory (32 t columns per 128-sponge tile, 512 columns per SM in all), so blocks of two kernels that share an SM compete for
it: small launches (one or two tiles per block, little shared memory) are the case where several blocks are resident on
one SM.  The launches go to separate streams back to back, many times; every result is compared with the oracle."""
import numpy as np
import pytest

from conftest import field_id, field_name, oracle_cfg

pytestmark = pytest.mark.gpu


def test_variant7_kernels_of_different_widths_share_the_sms(ref):
    import torch
    import sponge_b200 as S
    from oracle.c_oracle import COracle
    names = ["bn254_t3", "bn254_t5", "bn254_t2", "bls_t3_alpha5"]
    ctxs, orcs, ts = [], [], []
    try:
        for name in names:
            oc = oracle_cfg(ref, name)
            cfg = S.PoseidonConfig(oc.full_rounds, oc.partial_rounds, oc.alpha, oc.mds, oc.ark, oc.rate, oc.capacity,
                                   field_name(ref, oc))
            ctx = cfg.context()
            assert ctx.active_variant() == 7, ctx.describe()
            ctxs.append(ctx)
            orcs.append(COracle(oc, field_id(ref, oc)))
            ts.append(oc.rate + oc.capacity)
        sms = torch.cuda.get_device_properties(0).multi_processor_count
        # one tile per block on every SM twice over / two tiles / a ragged handful
        for n in (2 * sms * 128, sms * 256 + 77, 300):
            streams = [torch.cuda.Stream() for _ in ctxs]
            ins = [orc.synth(n * t, seed=0xCC0 + t) for orc, t in zip(orcs, ts)]
            exp = [orc.permute_batch(x.copy()) for orc, x in zip(orcs, ins)]
            bufs = [torch.from_numpy(x.copy().view(np.int64)).cuda() for x in ins]
            work = [b.clone() for b in bufs]
            for rep in range(12):
                for w, b in zip(work, bufs):
                    w.copy_(b)
                torch.cuda.synchronize()
                for ctx, w, s in zip(ctxs, work, streams):
                    ctx.permute_batch_dev(w.data_ptr(), n, stream=s.cuda_stream)
                torch.cuda.synchronize()
                for k, (w, e) in enumerate(zip(work, exp)):
                    assert np.array_equal(w.cpu().numpy().view(np.uint64), e), (names[k], n, rep)
    finally:
        for ctx in ctxs:
            ctx.close()
