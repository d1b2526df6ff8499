"""psd_merkle_root_noalloc: the Merkle entry point for hosts that may not allocate on the request
path (their arena owns all device memory, so a library-side cudaMalloc can fail and stalls every
stream).  It keeps the intermediate tree levels in a fixed 1.5 MiB array inside the library instead
of a cudaMalloc'd buffer.

Checked here, through the C ABI: its root is bit-identical (the tolerance every parity test in this
repository uses for integer work) to the one psd_merkle_root returns for the same leaves, on two
calls in a row that share the scratch array -- the second with a different, smaller tree, so
anything left behind by the first call would show up -- and sizes beyond the fixed maximum are
refused with a clear message.  This file sorts last on purpose: a run that stops at the first
failure has already executed every other test.  It passes on its own in a fresh process as well as
after the rest of the suite (the set-up of the scratch address happens on the first call either way).
"""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def test_noalloc_merkle_root_matches_allocating_entry_point():
    import torch
    import sponge_b200 as S
    from sponge_b200 import _lib as L
    from oracle import poseidon_ref as R
    from oracle.c_oracle import COracle

    lib = L.lib
    ark, mds = S.find_poseidon_ark_and_mds("bn254", 254, 2, 8, 57, 0)
    cfg = S.PoseidonConfig.new(8, 57, 5, mds, ark, 2, 1, "bn254")
    ctx = cfg.context(0)
    orc = COracle(R.bn254_config(2, 8, 57), 0)   # the CPU oracle: the check does not rest on psd_merkle_root alone
    try:
        dev = torch.device("cuda:0")
        stream = torch.cuda.current_stream(dev).cuda_stream
        max_leaves = 65536   # PSD_NOALLOC_MAX_LEAVES

        def roots_for(n_leaves, seed, repr_):
            leaves = torch.empty((n_leaves, 4), dtype=torch.int64, device=dev)
            ctx.synth_fill_dev(leaves.data_ptr(), n_leaves, seed, repr=repr_, stream=stream)
            want = torch.zeros((1, 4), dtype=torch.int64, device=dev)
            got = torch.zeros((1, 4), dtype=torch.int64, device=dev)
            # existing entry point (allocates its level buffers with cudaMalloc)
            L.check(lib.psd_merkle_root(ctx.handle, leaves.data_ptr(), n_leaves, want.data_ptr(), repr_, L.MEM_DEVICE,
                                        stream))
            # new entry point: same inputs, caller-owned output, no allocation inside
            L.check(lib.psd_merkle_root_noalloc(ctx.handle, leaves.data_ptr(), n_leaves, got.data_ptr(), repr_, stream))
            ctx.synchronize(stream)
            # the same leaves on the host (the synthetic generator is identical on both sides) through the oracle
            exp = orc.merkle_root(orc.synth(n_leaves, seed))
            if repr_ == L.REPR_CANONICAL:
                exp = orc.from_mont(exp.copy())
            assert np.array_equal(got.cpu().numpy().view(np.uint64).reshape(4), exp)
            return want.cpu().numpy().view(np.uint64), got.cpu().numpy().view(np.uint64)

        # call 1: the largest tree the fixed scratch array can hold, Montgomery leaves
        want1, got1 = roots_for(max_leaves, 0x5A1, L.REPR_MONTGOMERY)
        assert np.array_equal(got1, want1)
        # call 2, straight after, on the same shared array: a smaller tree, canonical leaves
        want2, got2 = roots_for(1 << 9, 0x5A2, L.REPR_CANONICAL)
        assert np.array_equal(got2, want2)
        assert not np.array_equal(got1, got2)
        # smallest non-trivial trees and the single-leaf case
        for n in (1, 2, 4):
            w, g = roots_for(n, 0x5A3 + n, L.REPR_MONTGOMERY)
            assert np.array_equal(g, w)

        # beyond the maximum: refused, with a message that names the limit
        leaves = torch.zeros((2 * max_leaves, 4), dtype=torch.int64, device=dev)
        out = torch.zeros((1, 4), dtype=torch.int64, device=dev)
        st = lib.psd_merkle_root_noalloc(ctx.handle, leaves.data_ptr(), 2 * max_leaves, out.data_ptr(),
                                         L.REPR_MONTGOMERY, stream)
        assert st == L.PSD_ERR_ARG
        assert b"PSD_NOALLOC_MAX_LEAVES" in lib.psd_last_error_string()
        # not a power of two / null pointers
        assert lib.psd_merkle_root_noalloc(ctx.handle, leaves.data_ptr(), 3, out.data_ptr(), L.REPR_MONTGOMERY,
                                           stream) == L.PSD_ERR_ARG
        assert lib.psd_merkle_root_noalloc(ctx.handle, None, 4, out.data_ptr(), L.REPR_MONTGOMERY, stream) == L.PSD_ERR_ARG
    finally:
        ctx.close()
