"""BASELINE.json configurations at their full sizes.  C1 is compared element for element with the
oracle; C2/C3/C4 are too large for a full CPU comparison, so they are checked through sampled
items against the oracle (first/last 2^12 items, every 65537th item, shard edges) plus
size-independent properties (sharded == unsharded Merkle root, the decomposition the multi-GPU mode
uses).  Inputs are generated on the device by psd_synth_fill, whose equality with the oracle's
generator is itself a test (test_gpu_parity.py::test_synth_fill_matches_oracle)."""
import numpy as np
import pytest

from conftest import field_id, field_name, oracle_cfg

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def S():
    import sponge_b200
    return sponge_b200


def make(S, ref, name):
    from oracle.c_oracle import COracle
    c = oracle_cfg(ref, name)
    cfg = S.PoseidonConfig(c.full_rounds, c.partial_rounds, c.alpha, c.mds, c.ark, c.rate, c.capacity, field_name(ref, c))
    return c, COracle(c, field_id(ref, c)), cfg.context()


def sample_indices(n):
    idx = np.concatenate([np.arange(0, 4096), np.arange(n - 4096, n), np.arange(0, n, 65537),
                          np.arange(n // 2 - 64, n // 2 + 64), np.arange(n // 8 - 16, n // 8 + 16)])
    return np.unique(idx[(idx >= 0) & (idx < n)])


@pytest.mark.parametrize("name", ["bls_fixture", "bls_default_r2"])
def test_c1_full_compare(S, ref, name):
    """configs[0]: 2^16 independent BLS12-381 sponges absorb 2 elements, squeeze 1 -- with the literal
    test fixture (src/poseidon/tests.rs:120-820) and with the KAT-pinned default set."""
    _, orc, ctx = make(S, ref, name)
    try:
        n = 1 << 16
        x = orc.synth(n * 2, seed=0xC1)
        exp = orc.absorb_squeeze_batch(x, n, 2, 1)
        assert np.array_equal(ctx.absorb_squeeze_batch(x, n, 2, 1), exp)
        xc = orc.from_mont(x.copy())
        got_c = ctx.absorb_squeeze_batch(xc, n, 2, 1, repr=S.REPR_CANONICAL)
        assert np.array_equal(got_c, orc.from_mont(exp.copy()))
    finally:
        ctx.close()


def test_c2_full_size_sampled(S, ref):
    """configs[1]: 2^24 independent BN254 t=3 permutations, device resident, in place."""
    import torch
    _, orc, ctx = make(S, ref, "bn254_t3")
    try:
        n, t = 1 << 24, 3
        buf = torch.empty((n * t, 4), dtype=torch.int64, device="cuda:0")
        ctx.synth_fill_dev(buf.data_ptr(), n * t, 0xC2)
        ctx.permute_batch_dev(buf.data_ptr(), n)
        ctx.synchronize()
        idx = sample_indices(n)
        got = buf.view(n, t * 4)[torch.from_numpy(idx).to("cuda:0")].cpu().numpy().view(np.uint64)
        inp = np.concatenate([orc.synth(t, 0xC2, first=int(i) * t) for i in idx])
        exp = orc.permute_batch(inp).reshape(len(idx), t * 4)
        assert np.array_equal(got, exp)
    finally:
        ctx.close()


def test_c3_full_size_sampled(S, ref):
    """configs[2]: 2^24 BN254 t=5 absorb-4 / squeeze-1 hashes."""
    import torch
    _, orc, ctx = make(S, ref, "bn254_t5")
    try:
        n = 1 << 24
        x = torch.empty((n * 4, 4), dtype=torch.int64, device="cuda:0")
        out = torch.empty((n, 4), dtype=torch.int64, device="cuda:0")
        ctx.synth_fill_dev(x.data_ptr(), n * 4, 0xC3)
        ctx.absorb_squeeze_batch_dev(x.data_ptr(), n, 4, 1, out.data_ptr())
        ctx.synchronize()
        idx = sample_indices(n)
        got = out[torch.from_numpy(idx).to("cuda:0")].cpu().numpy().view(np.uint64)
        inp = np.concatenate([orc.synth(4, 0xC3, first=int(i) * 4) for i in idx])
        exp = orc.absorb_squeeze_batch(inp, len(idx), 4, 1).reshape(len(idx), 4)
        assert np.array_equal(got, exp)
    finally:
        ctx.close()


def test_c4_merkle_full_size_properties(S, ref):
    """configs[3]: a 2^26-leaf tree.  (i) the root of the whole tree equals the root over the 8 subtree
    roots -- exactly the per-GPU decomposition of the multi-GPU mode; (ii) sampled 2^10-leaf subtrees
    equal the oracle's roots; (iii) changing one leaf changes the root."""
    import torch
    _, orc, ctx = make(S, ref, "bn254_t3")
    try:
        n = 1 << 26
        leaves = torch.empty((n, 4), dtype=torch.int64, device="cuda:0")
        ctx.synth_fill_dev(leaves.data_ptr(), n, 0xC4)
        root = torch.empty((1, 4), dtype=torch.int64, device="cuda:0")
        ctx.merkle_root_dev(leaves.data_ptr(), n, root.data_ptr())
        sub = torch.empty((8, 4), dtype=torch.int64, device="cuda:0")
        per = n // 8
        for g in range(8):
            ctx.merkle_root_dev(leaves.data_ptr() + g * per * 32, per, sub.data_ptr() + g * 32)
        top = torch.empty((1, 4), dtype=torch.int64, device="cuda:0")
        ctx.merkle_root_dev(sub.data_ptr(), 8, top.data_ptr())
        ctx.synchronize()
        assert torch.equal(root, top)
        small = torch.empty((1, 4), dtype=torch.int64, device="cuda:0")
        for start in (0, per - 1024, 5 * per + 77 * 1024, n - 1024):
            ctx.merkle_root_dev(leaves.data_ptr() + start * 32, 1024, small.data_ptr())
            ctx.synchronize()
            exp = orc.merkle_root(orc.synth(1024, 0xC4, first=start))
            assert np.array_equal(small.cpu().numpy().view(np.uint64).reshape(4), exp)
        leaves[12345678, 0] += 1          # still a valid Montgomery residue with overwhelming probability
        root2 = torch.empty((1, 4), dtype=torch.int64, device="cuda:0")
        ctx.merkle_root_dev(leaves.data_ptr(), n, root2.data_ptr())
        ctx.synchronize()
        assert not torch.equal(root, root2)
    finally:
        ctx.close()


def test_c5_shape_long_absorb_in_chunks(S, ref):
    """configs[4] shape at reduced width (2^12 sponges x 1024 elements, squeeze_bytes(32)): absorbing in
    16 chunks of 64 through the lock-step batch equals the one-shot call and the oracle."""
    _, orc, ctx = make(S, ref, "bls_fixture")
    try:
        n, k, chunk = 1 << 12, 1024, 64
        x = orc.synth(n * k, seed=0xC5)
        exp = orc.absorb_squeeze_bytes_batch(x, n, k, 32)
        assert np.array_equal(ctx.absorb_squeeze_bytes_batch(x, n, k, 32), exp)
        b = S.SpongeBatch(ctx, n)
        xs = x.reshape(n, k, 4)
        for c in range(k // chunk):
            b.absorb(np.ascontiguousarray(xs[:, c * chunk:(c + 1) * chunk]), chunk)
        assert np.array_equal(b.squeeze_bytes(32), exp)
        b.free()
    finally:
        ctx.close()


@pytest.mark.parametrize("name,absorb", [("bn254_t3", 0), ("bn254_t5", 4)])
def test_fast_variants_equal_dense_on_every_element_of_the_full_batch(S, ref, name, absorb):
    """The fast kernels (variant 6: tensor-core linear layer; variant 5: 52-bit limbs on the FP64 pipe, lazily
    reduced lanes) against the dense reference-order kernel (variant 0, itself oracle-tested) on ALL 2^24 items
    of C2 / C3, compared on the device -- not a sample: a carry corner that fires once in 2^24 inputs shows up
    here."""
    import torch
    _, orc, ctx = make(S, ref, name)
    try:
        n = 1 << 24
        t = 3 if absorb == 0 else 5
        width = t if absorb == 0 else absorb
        x = torch.empty((n * width, 4), dtype=torch.int64, device="cuda:0")
        ctx.synth_fill_dev(x.data_ptr(), n * width, 0x5D0)
        outs = []
        for variant in (7, 6, 0):
            ctx.set_variant(variant)
            assert ctx.active_variant() == variant
            if absorb == 0:
                y = x.clone()
                ctx.permute_batch_dev(y.data_ptr(), n)
            else:
                y = torch.empty((n, 4), dtype=torch.int64, device="cuda:0")
                ctx.absorb_squeeze_batch_dev(x.data_ptr(), n, absorb, 1, y.data_ptr())
            ctx.synchronize()
            outs.append(y)
        assert torch.equal(outs[0], outs[2])
        assert torch.equal(outs[1], outs[2])
        # anchor them to the oracle on a few items
        idx = np.array([0, 1, n // 3, n - 1])
        got = outs[0].view(n, -1)[torch.from_numpy(idx).to("cuda:0")].cpu().numpy().view(np.uint64)
        inp = np.concatenate([orc.synth(width, 0x5D0, first=int(i) * width) for i in idx])
        exp = orc.permute_batch(inp) if absorb == 0 else orc.absorb_squeeze_batch(inp, len(idx), absorb, 1)
        assert np.array_equal(got.reshape(-1), exp.reshape(-1))
    finally:
        ctx.close()


def test_c5_full_width_sampled(S, ref):
    """configs[4] at its full width: 2^20 BLS12-381 sponges each absorb 1024 elements, then squeeze_bytes(32).
    The 32 GiB of inputs are never materialised: every sponge's inputs are the counter-based synthetic stream
    regenerated on the device chunk by chunk (64 elements per sponge and call, distinct elements throughout),
    absorbed through the lock-step batch; a 2^12-item sample including both ends and the 8-way shard edges is
    recomputed by the oracle from the same stream."""
    import torch
    _, orc, ctx = make(S, ref, "bls_fixture")
    try:
        n, k, chunk = 1 << 20, 1024, 64
        buf = torch.empty((n * chunk, 4), dtype=torch.int64, device="cuda:0")
        outb = torch.empty((n, 32), dtype=torch.uint8, device="cuda:0")
        b = S.SpongeBatch(ctx, n)
        s = torch.cuda.current_stream().cuda_stream
        b.set_stream(s)
        # chunk c of sponge i = stream elements [(c*n + i)*chunk, (c*n + i + 1)*chunk)
        for c in range(k // chunk):
            ctx.synth_fill_dev(buf.data_ptr(), n * chunk, 0xC5, first_index=c * n * chunk, stream=s)
            b.absorb_dev(buf.data_ptr(), chunk)
        b.squeeze_bytes_dev(32, outb.data_ptr())
        ctx.synchronize(s)
        edges = [g * (n // 8) + d for g in range(1, 8) for d in (-1, 0)]
        idx = np.unique(np.concatenate([np.arange(0, 1024), np.arange(n - 1024, n), np.arange(0, n, 511),
                                        np.array(edges)]))[:4096]
        got = outb[torch.from_numpy(idx).to("cuda:0")].cpu().numpy()
        inp = np.concatenate([orc.synth(chunk, 0xC5, first=(c * n + int(i)) * chunk)
                              for i in idx for c in range(k // chunk)])
        exp = orc.absorb_squeeze_bytes_batch(inp, len(idx), k, 32)
        assert np.array_equal(got, exp)
        b.free()
    finally:
        ctx.close()
