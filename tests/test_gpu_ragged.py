"""Ragged batches: every sponge absorbs its own number of elements (psd_absorb_squeeze_ragged_batch), as the
reference's own behavioural test does with lists of differing lengths (src/poseidon/tests.rs:45-69).  Checked
against the oracle, which hashes the sponges bucket by bucket (one lock-step call per distinct length)."""
import numpy as np
import pytest

from conftest import field_id, field_name, oracle_cfg

pytestmark = pytest.mark.gpu


def make(ref, name):
    import sponge_b200 as S
    from oracle.c_oracle import COracle
    c = oracle_cfg(ref, name)
    cfg = S.PoseidonConfig(c.full_rounds, c.partial_rounds, c.alpha, c.mds, c.ark, c.rate, c.capacity, field_name(ref, c))
    return S, COracle(c, field_id(ref, c)), cfg.context()


def oracle_ragged(orc, x, offsets, n_squeeze):
    n = len(offsets) - 1
    out = np.zeros((n, n_squeeze, 4), dtype=np.uint64)
    lens = np.diff(offsets.astype(np.int64))
    for k in np.unique(lens):
        idx = np.nonzero(lens == k)[0]
        if k == 0:
            inp = np.zeros((0, 4), dtype=np.uint64)
        else:
            inp = np.concatenate([x[int(offsets[i]):int(offsets[i + 1])] for i in idx])
        out[idx] = orc.absorb_squeeze_batch(np.ascontiguousarray(inp), len(idx), int(k), n_squeeze)
    return out


@pytest.mark.parametrize("name", ["bn254_t3", "bn254_t5", "bls_fixture", "bn254_t4_cap2", "bls_default_r8", "bn254_t2"])
def test_ragged_absorb_matches_oracle(ref, name):
    S, orc, ctx = make(ref, name)
    try:
        rng = np.random.default_rng(0x7A6)
        n = 2500
        lens = rng.integers(0, 19, size=n)
        lens[:40] = np.arange(40) % 13              # every small length inside one warp
        lens[100:132] = 7                           # one uniform warp
        lens[200] = 300                             # one long sponge among short ones
        offsets = np.concatenate([[0], np.cumsum(lens)]).astype(np.uint64)
        x = orc.synth(int(offsets[-1]), seed=0x7A7)
        for n_squeeze in (1, 3):
            exp = oracle_ragged(orc, x, offsets, n_squeeze)
            got = ctx.absorb_squeeze_ragged_batch(x, offsets, n_squeeze)
            assert np.array_equal(got, exp), (name, n_squeeze, ctx.describe())
        # canonical representation in and out
        xc = orc.from_mont(x.copy())
        got_c = ctx.absorb_squeeze_ragged_batch(xc, offsets, 1, repr=S.REPR_CANONICAL)
        assert np.array_equal(got_c, orc.from_mont(oracle_ragged(orc, x, offsets, 1).reshape(-1, 4)).reshape(n, 1, 4))
        # equal lengths: the ragged call equals the lock-step call
        eq = np.arange(0, 5 * 64 + 1, 5, dtype=np.uint64)
        xe = orc.synth(5 * 64, seed=0x7A8)
        assert np.array_equal(ctx.absorb_squeeze_ragged_batch(xe, eq, 2), ctx.absorb_squeeze_batch(xe, 64, 5, 2))
    finally:
        ctx.close()


def test_ragged_large_and_argument_errors(ref):
    import ctypes as C
    S, orc, ctx = make(ref, "bn254_t3")
    from sponge_b200 import _lib as L
    try:
        # large enough for several staging chunks (more than 2^17 sponges) with mixed lengths
        rng = np.random.default_rng(5)
        n = (1 << 17) * 2 + 77
        lens = rng.integers(1, 6, size=n)
        offsets = np.concatenate([[0], np.cumsum(lens)]).astype(np.uint64)
        x = orc.synth(int(offsets[-1]), seed=0x7A9)
        got = ctx.absorb_squeeze_ragged_batch(x, offsets, 1)
        idx = np.concatenate([np.arange(0, 300), np.arange((1 << 17) - 150, (1 << 17) + 150), np.arange(n - 300, n),
                              rng.integers(0, n, size=600)])
        sub_off = np.concatenate([[0], np.cumsum(lens[idx])]).astype(np.uint64)
        sub_x = np.concatenate([x[int(offsets[i]):int(offsets[i + 1])] for i in idx])
        assert np.array_equal(got[idx], oracle_ragged(orc, sub_x, sub_off, 1))
        # errors
        out = np.zeros((2, 1, 4), dtype=np.uint64)
        bad = np.array([0, 3, 2], dtype=np.uint64)                    # decreasing
        st = L.lib.psd_absorb_squeeze_ragged_batch(ctx.handle, x.ctypes.data, bad.ctypes.data, 2, 1, out.ctypes.data,
                                                   L.REPR_MONTGOMERY, L.MEM_HOST, None)
        assert st == L.PSD_ERR_ARG
        bad0 = np.array([1, 2, 3], dtype=np.uint64)                   # does not start at 0
        st = L.lib.psd_absorb_squeeze_ragged_batch(ctx.handle, x.ctypes.data, bad0.ctypes.data, 2, 1, out.ctypes.data,
                                                   L.REPR_MONTGOMERY, L.MEM_HOST, None)
        assert st == L.PSD_ERR_ARG
        # a non-canonical element is reported
        xc = orc.from_mont(x[:8].copy())
        xc[3] = np.array([0xffffffffffffffff] * 4, dtype=np.uint64)
        off = np.array([0, 2, 8], dtype=np.uint64)
        with pytest.raises(S.PoseidonB200Error) as ei:
            ctx.absorb_squeeze_ragged_batch(xc, off, 1, repr=S.REPR_CANONICAL)
        assert ei.value.status == L.PSD_ERR_NONCANONICAL
    finally:
        ctx.close()
