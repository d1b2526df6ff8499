"""Several GPUs in ONE process through psd_permute_batch_multi / psd_absorb_squeeze_batch_multi /
psd_merkle_root_multi, against the oracle.  Skipped below two devices; on a larger box every power-of-two
device count up to the number present is exercised.  Shard edges get explicit attention: the items on both
sides of every boundary between two devices' ranges are compared with the oracle one by one."""
import numpy as np
import pytest

from conftest import field_id, field_name, oracle_cfg

pytestmark = pytest.mark.gpu


def _device_counts():
    import sponge_b200._lib as L
    n = L.lib.psd_device_count()
    return [g for g in (2, 4, 8) if g <= n]


@pytest.fixture(scope="module")
def env(ref):
    import sponge_b200 as S
    from oracle.c_oracle import COracle
    counts = _device_counts()
    if not counts:
        pytest.skip("needs at least two CUDA devices")
    oc = oracle_cfg(ref, "bn254_t3")
    cfg = S.PoseidonConfig(oc.full_rounds, oc.partial_rounds, oc.alpha, oc.mds, oc.ark, oc.rate, oc.capacity,
                           field_name(ref, oc))
    return S, cfg, COracle(oc, field_id(ref, oc)), counts


def test_sharded_permute_and_hash_match_oracle_at_every_shard_edge(env):
    S, cfg, orc, counts = env
    for g in counts:
        mc = S.MultiContext(cfg, list(range(g)))
        try:
            for n in (g * 4096 + 3, 5, 1 << 18):          # ragged, fewer items than devices, multi-chunk
                st = orc.synth(n * 3, seed=0x3A + g)
                got = mc.permute_batch(st.copy())
                if n <= 1 << 15:
                    assert np.array_equal(got, orc.permute_batch(st.copy()))
                edges = sorted({min(max(e + d, 0), n - 1) for r in range(g) for e in S.shard_range(n, r, g) for d in (-1, 0)})
                idx = np.array(edges + [0, n // 2, n - 1])
                exp = orc.permute_batch(st.reshape(n, 12)[idx].copy().reshape(-1, 4)).reshape(len(idx), 12)
                assert np.array_equal(got.reshape(n, 12)[idx], exp), (g, n)
            # canonical + SoA through the sharded path
            n = g * 1000 + 1
            st = orc.synth(n * 3, seed=0x3B)
            exp = orc.permute_batch(st.copy())
            soa = np.ascontiguousarray(st.reshape(n, 3, 2, 2).transpose(1, 2, 0, 3))
            got = mc.permute_batch(soa, layout=S.LAYOUT_SOA)
            assert np.array_equal(np.ascontiguousarray(got.reshape(3, 2, n, 2).transpose(2, 0, 1, 3)).reshape(-1, 4), exp)
            got_c = mc.permute_batch(orc.from_mont(st.copy()), repr=S.REPR_CANONICAL)
            assert np.array_equal(got_c, orc.from_mont(exp.copy()))
            # absorb-5 / squeeze-2 hashes
            x = orc.synth(n * 5, seed=0x3C)
            assert np.array_equal(mc.absorb_squeeze_batch(x, n, 5, 2), orc.absorb_squeeze_batch(x, n, 5, 2))
            # multi == single device
            one = mc.ctxs[0].permute_batch(st.copy())
            assert np.array_equal(one, exp)
        finally:
            mc.close()


def test_multi_device_merkle_root_equals_single_device_and_oracle(env):
    import torch
    S, cfg, orc, counts = env
    for g in counts:
        mc = S.MultiContext(cfg, list(range(g)))
        try:
            for per in (1 << 14, 2, 1):
                leaves = orc.synth(per * g, seed=0x4D + g)
                want = orc.merkle_root(leaves)
                assert np.array_equal(mc.merkle_root(leaves), want), (g, per)
                assert np.array_equal(mc.ctxs[0].merkle_root(leaves), want)
                canon = orc.from_mont(leaves.copy())
                assert np.array_equal(mc.merkle_root(canon, repr=S.REPR_CANONICAL), orc.from_mont(want.copy()))
            # device-resident leaves, one buffer per GPU, produced on each device's current torch stream
            per = 1 << 16
            bufs, streams = [], []
            for d in range(g):
                with torch.cuda.device(d):
                    b = torch.empty((per, 4), dtype=torch.int64, device=f"cuda:{d}")
                    s = torch.cuda.current_stream(d).cuda_stream
                    mc.ctxs[d].synth_fill_dev(b.data_ptr(), per, 0x4E, first_index=d * per, stream=s)
                    bufs.append(b)
                    streams.append(s)
            root = torch.zeros((1, 4), dtype=torch.int64, device="cuda:0")
            mc.merkle_root_dev([b.data_ptr() for b in bufs], per, root.data_ptr(), streams=streams)
            want = orc.merkle_root(orc.synth(per * g, seed=0x4E))
            assert np.array_equal(root.cpu().numpy().view(np.uint64).reshape(4), want)
            # a second call on the same contexts (buffers and events are reused)
            mc.merkle_root_dev([b.data_ptr() for b in bufs], per, root.data_ptr(), streams=streams)
            assert np.array_equal(root.cpu().numpy().view(np.uint64).reshape(4), want)
        finally:
            mc.close()


def test_multi_argument_errors(env):
    import ctypes as C
    import sponge_b200._lib as L
    S, cfg, orc, counts = env
    mc = S.MultiContext(cfg, [0, 1])
    try:
        lib = L.lib
        arr = (C.c_void_p * 2)(mc.ctxs[0].handle, mc.ctxs[0].handle)       # the same context twice
        st = orc.synth(6, seed=1)
        assert lib.psd_permute_batch_multi(arr, 2, st.ctypes.data, 2, L.LAYOUT_AOS, L.REPR_MONTGOMERY) == L.PSD_ERR_ARG
        other = S.PoseidonConfig(cfg.full_rounds, cfg.partial_rounds, cfg.alpha, cfg.mds,
                                 [list(reversed(r)) for r in cfg.ark], cfg.rate, cfg.capacity, cfg.field)
        oc = S.PoseidonContext(other, 1)
        try:
            arr = (C.c_void_p * 2)(mc.ctxs[0].handle, oc.handle)           # different parameter sets
            assert lib.psd_permute_batch_multi(arr, 2, st.ctypes.data, 2, L.LAYOUT_AOS, L.REPR_MONTGOMERY) == L.PSD_ERR_ARG
            assert b"different parameter set" in lib.psd_last_error_string()
        finally:
            oc.close()
        leaves = orc.synth(6, seed=2)
        with pytest.raises(S.PoseidonB200Error):
            mc.merkle_root(leaves)                                           # 3 leaves per device
    finally:
        mc.close()
