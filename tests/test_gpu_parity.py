"""GPU parity: the CUDA path, called through the C ABI (sponge_b200 -> libposeidon_b200.so),
against the CPU oracle on the same seeded inputs.  Bit-exact (integer work): every comparison is
array equality on the 32-byte limb images, in both canonical and Montgomery representations."""
import numpy as np
import pytest

from conftest import field_id, field_name, oracle_cfg, to_ints, to_limbs, load_golden

pytestmark = pytest.mark.gpu

CFG_NAMES = ["bn254_t3", "bn254_t5", "bls_default_r2", "bls_fixture", "bls_default_r3", "bls_default_r2_w",
             "bn254_t2", "bn254_t4_cap2", "bn254_odd_rf", "bls_default_r5", "bls_default_r8",
             "bn254_full_only", "bn254_rf1", "bls_t3_alpha5"]


@pytest.fixture(scope="module")
def S():
    import sponge_b200
    return sponge_b200


class Pair:
    """One parameter set on both sides: GPU context + C oracle."""

    def __init__(self, S, ref, name):
        from oracle.c_oracle import COracle
        self.ocfg = oracle_cfg(ref, name)
        self.orc = COracle(self.ocfg, field_id(ref, self.ocfg))
        c = self.ocfg
        self.cfg = S.PoseidonConfig(c.full_rounds, c.partial_rounds, c.alpha, c.mds, c.ark, c.rate, c.capacity,
                                    field_name(ref, c))
        self.ctx = self.cfg.context()
        self.t = c.rate + c.capacity
        self.p = c.p

    def close(self):
        self.ctx.close()


@pytest.fixture(scope="module")
def pairs(S, ref):
    cache = {}

    def get(name):
        if name not in cache:
            while len(cache) >= 2:   # the 62 KiB constant bank holds only a few parameter sets at once
                old = next(iter(cache))
                cache.pop(old).close()
            cache[name] = Pair(S, ref, name)
        return cache[name]

    yield get
    for p in cache.values():
        p.close()


def soa(a, n, w):
    """(n, w, 4) AoS -> SoA image [w][2][n][2]."""
    return np.ascontiguousarray(a.reshape(n, w, 2, 2).transpose(1, 2, 0, 3))


def unsoa(a, n, w):
    return np.ascontiguousarray(a.reshape(w, 2, n, 2).transpose(2, 0, 1, 3)).reshape(n, w, 4)


@pytest.mark.parametrize("name", CFG_NAMES)
def test_permute_batch_matches_oracle(S, pairs, name):
    pr = pairs(name)
    n = 3000 if pr.t <= 5 else 600
    st = pr.orc.synth(n * pr.t, seed=0xA1, montgomery=True)
    # edge states: all zero, all p-1
    st[: pr.t] = 0
    st[pr.t: 2 * pr.t] = pr.orc.to_mont(to_limbs([pr.p - 1] * pr.t))
    exp = pr.orc.permute_batch(st.copy())
    got = pr.ctx.permute_batch(st.copy())
    assert np.array_equal(got, exp)
    # canonical representation in and out
    canon_in = pr.orc.from_mont(st.copy())
    got_c = pr.ctx.permute_batch(canon_in.copy(), repr=S.REPR_CANONICAL)
    assert np.array_equal(got_c, pr.orc.from_mont(exp.copy()))
    # SoA layout
    got_s = pr.ctx.permute_batch(soa(st, n, pr.t), layout=S.LAYOUT_SOA)
    assert np.array_equal(unsoa(got_s, n, pr.t).reshape(-1, 4), exp)


@pytest.mark.parametrize("name", ["bn254_t3", "bls_default_r2", "bn254_t5", "bls_fixture", "bn254_t4_cap2"])
def test_permute_dense_variant_equals_optimised(S, pairs, name):
    pr = pairs(name)
    st = pr.orc.synth(2048 * pr.t, seed=0xA2)
    pr.ctx.set_variant(0)
    dense = pr.ctx.permute_batch(st.copy())
    pr.ctx.set_variant(1)
    opt = pr.ctx.permute_batch(st.copy())
    pr.ctx.set_variant(2)
    compact = pr.ctx.permute_batch(st.copy())
    pr.ctx.set_variant(3)          # digit tables (falls back to 2 where the context has none); the default
    ctab = pr.ctx.permute_batch(st.copy())
    assert np.array_equal(dense, opt)
    assert np.array_equal(dense, compact)
    assert np.array_equal(dense, ctab)
    assert np.array_equal(dense, pr.orc.permute_batch(st.copy()))


def test_circomlib_vector_bn254(S, pairs):
    """External anchor for BN254 (not in the reference): circomlib poseidon([1,2])."""
    pr = pairs("bn254_t3")
    st = to_limbs([0, 1, 2])
    out = to_ints(pr.ctx.permute_batch(st, repr=S.REPR_CANONICAL))
    assert out[0] == 0x115cc0f5e7d690413df64c6b9662e9cf2a3617f2743245519e19607a4417189a


@pytest.mark.parametrize("name", CFG_NAMES)
@pytest.mark.parametrize("n_absorb,n_squeeze", [(2, 1), (1, 1), (0, 1), (4, 1), (7, 5), (3, 3), (16, 0), (5, 9)])
def test_absorb_squeeze_batch(S, pairs, name, n_absorb, n_squeeze):
    pr = pairs(name)
    n = 700 if pr.t <= 5 else 200
    x = pr.orc.synth(max(n * n_absorb, 1), seed=0xB0 + n_absorb)[: n * n_absorb]
    exp = pr.orc.absorb_squeeze_batch(x, n, n_absorb, n_squeeze)
    got = pr.ctx.absorb_squeeze_batch(x, n, n_absorb, n_squeeze)
    assert np.array_equal(got, exp)
    if n_absorb and n_squeeze:
        got_s = pr.ctx.absorb_squeeze_batch(soa(x, n, n_absorb), n, n_absorb, n_squeeze, layout=S.LAYOUT_SOA)
        assert np.array_equal(unsoa(got_s, n, n_squeeze), exp)
        xc = pr.orc.from_mont(x.copy())
        got_c = pr.ctx.absorb_squeeze_batch(xc, n, n_absorb, n_squeeze, repr=S.REPR_CANONICAL)
        assert np.array_equal(got_c, pr.orc.from_mont(exp.copy()))


@pytest.mark.parametrize("name", ["bls_fixture", "bn254_t3", "bn254_t5"])
@pytest.mark.parametrize("n_absorb,n_bytes", [(2, 32), (5, 31), (3, 1), (9, 100), (4, 0), (64, 32)])
def test_absorb_squeeze_bytes_batch(S, pairs, name, n_absorb, n_bytes):
    pr = pairs(name)
    n = 300
    x = pr.orc.synth(n * n_absorb, seed=0xC0 + n_bytes)
    exp = pr.orc.absorb_squeeze_bytes_batch(x, n, n_absorb, n_bytes)
    got = pr.ctx.absorb_squeeze_bytes_batch(x, n, n_absorb, n_bytes)
    assert np.array_equal(got, exp)


def test_config5_long_absorb_golden(S, pairs):
    """BASELINE.json config 5 shape: 1024-element absorb then squeeze_bytes(32); golden from the
    KAT-gated Python oracle (tests/golden/derived_vectors.json)."""
    pr = pairs("bls_fixture")
    g = load_golden("derived_vectors.json")["bls_fixture_absorb_0_1023_squeeze_bytes_32"]
    x = to_limbs(range(1024))
    got = pr.ctx.absorb_squeeze_bytes_batch(x, 1, 1024, 32, repr=S.REPR_CANONICAL)
    assert bytes(got[0]).hex() == g
    # and as a batch, against the C oracle
    n = 64
    xs = pr.orc.synth(n * 1024, seed=0xC5)
    assert np.array_equal(pr.ctx.absorb_squeeze_bytes_batch(xs, n, 1024, 32),
                          pr.orc.absorb_squeeze_bytes_batch(xs, n, 1024, 32))


@pytest.mark.parametrize("name", ["bn254_t3", "bls_fixture", "bn254_t5"])
@pytest.mark.parametrize("n_leaves", [1, 2, 8, 1024, 1 << 14])
def test_merkle_root(S, pairs, name, n_leaves):
    pr = pairs(name)
    leaves = pr.orc.synth(n_leaves, seed=0xD0)
    exp = pr.orc.merkle_root(leaves)
    assert np.array_equal(pr.ctx.merkle_root(leaves), exp)
    canon = pr.orc.from_mont(leaves.copy())
    got_c = pr.ctx.merkle_root(canon, repr=S.REPR_CANONICAL)
    assert np.array_equal(got_c, pr.orc.from_mont(exp.copy()))


def test_merkle_golden(S, pairs):
    g = load_golden("derived_vectors.json")
    pr = pairs("bn254_t3")
    r8 = to_ints(pr.ctx.merkle_root(to_limbs(range(8)), repr=S.REPR_CANONICAL))[0]
    assert r8 == int(g["bn254_t3_merkle_leaves_0_7"], 16)
    r1024 = to_ints(pr.ctx.merkle_root(to_limbs(range(1, 1025)), repr=S.REPR_CANONICAL))[0]
    assert r1024 == int(g["bn254_t3_merkle_leaves_1_1024"], 16)


@pytest.mark.parametrize("repr_name", ["mont", "canon"])
def test_merkle_tree_and_paths(S, pairs, repr_name):
    """psd_merkle_tree returns every internal node; each level equals the oracle's, the last node is
    psd_merkle_root's root, and an authentication path recomputes the root through compress_pairs."""
    pr = pairs("bn254_t3")
    n = 1 << 10
    leaves_m = pr.orc.synth(n, seed=0xD7)
    canon = repr_name == "canon"
    leaves = pr.orc.from_mont(leaves_m.copy()) if canon else leaves_m
    rep = S.REPR_CANONICAL if canon else S.REPR_MONTGOMERY
    nodes = pr.ctx.merkle_tree(leaves, repr=rep)
    assert nodes.shape == (n - 1, 4)
    level, off = leaves_m, 0
    while level.shape[0] > 1:
        nxt = pr.orc.absorb_squeeze_batch(np.ascontiguousarray(level), level.shape[0] // 2, 2, 1).reshape(-1, 4)
        want = pr.orc.from_mont(nxt.copy()) if canon else nxt
        assert np.array_equal(nodes[off:off + nxt.shape[0]], want)
        off += nxt.shape[0]
        level = nxt
    assert np.array_equal(nodes[-1], pr.ctx.merkle_root(leaves, repr=rep))
    for idx in (0, 1, 517, n - 1):
        path = pr.ctx.merkle_path(leaves, nodes, idx)
        assert path.shape == (10, 4)
        cur, pos = leaves[idx], idx
        for sib in path:
            pair = np.stack([cur, sib]) if pos % 2 == 0 else np.stack([sib, cur])
            cur = pr.ctx.compress_pairs(np.ascontiguousarray(pair), repr=rep)[0]
            pos //= 2
        assert np.array_equal(cur, nodes[-1])


def test_compress_pairs(S, pairs):
    pr = pairs("bn254_t3")
    x = pr.orc.synth(2 * 999, seed=0xD1)
    assert np.array_equal(pr.ctx.compress_pairs(x), pr.orc.absorb_squeeze_batch(x, 999, 2, 1).reshape(-1, 4))


def test_montgomery_round_trip_and_constants(S, pairs):
    for name in ("bn254_t3", "bls_fixture"):
        pr = pairs(name)
        x = pr.orc.synth(5000, seed=0xE0, montgomery=False)
        m = pr.ctx.to_montgomery(x.copy())
        assert np.array_equal(m, pr.orc.to_mont(x.copy()))
        assert np.array_equal(pr.ctx.from_montgomery(m.copy()), x)
        one = pr.ctx.to_montgomery(to_limbs([1]))
        assert to_ints(one)[0] == (1 << 256) % pr.p


def test_noncanonical_input_rejected(S, pairs):
    pr = pairs("bn254_t3")
    bad = to_limbs([pr.p, 1, 2])
    with pytest.raises(S.PoseidonB200Error) as e:
        pr.ctx.permute_batch(bad, repr=S.REPR_CANONICAL)
    assert e.value.status == 3   # PSD_ERR_NONCANONICAL
    # the flag is cleared: the next call succeeds
    ok = pr.ctx.permute_batch(to_limbs([0, 1, 2]), repr=S.REPR_CANONICAL)
    assert to_ints(ok)[0] == 0x115cc0f5e7d690413df64c6b9662e9cf2a3617f2743245519e19607a4417189a


def test_synth_fill_matches_oracle(S, pairs):
    import torch
    for name in ("bn254_t3", "bls_fixture"):
        pr = pairs(name)
        n = 10000
        buf = torch.empty((n, 4), dtype=torch.int64, device="cuda:0")
        pr.ctx.synth_fill_dev(buf.data_ptr(), n, 0xC2, 77, repr=S.REPR_MONTGOMERY)
        pr.ctx.synchronize()
        got = buf.cpu().numpy().view(np.uint64)
        assert np.array_equal(got, pr.orc.synth(n, 0xC2, 77, montgomery=True))


def test_device_pointer_path_with_torch_stream(S, pairs):
    import torch
    pr = pairs("bn254_t3")
    n = 1 << 15
    st = pr.orc.synth(n * 3, seed=0xF1)
    d = torch.from_numpy(st.view(np.int64)).to("cuda:0")
    stream = torch.cuda.Stream()
    with torch.cuda.stream(stream):
        pr.ctx.permute_batch_dev(d.data_ptr(), n, stream=stream.cuda_stream)
    stream.synchronize()
    assert np.array_equal(d.cpu().numpy().view(np.uint64), pr.orc.permute_batch(st.copy()))


def test_large_batch_properties(S, pairs):
    """Size-independent checks at a larger size through the staged host path (the library splits a
    host batch into chunks of 2^16 sponges over three streams): sampled states against the oracle,
    including both sides of chunk boundaries, and the permutation is a bijection on the sample
    (distinct inputs -> distinct outputs)."""
    pr = pairs("bn254_t3")
    n = (1 << 20) + 12345
    st = pr.orc.synth(n * 3, seed=0xF2)
    got = pr.ctx.permute_batch(st.copy()).reshape(n, 12)
    idx = np.unique(np.concatenate([np.arange(0, 4096), np.arange(n - 4096, n), np.arange(0, n, 257),
                                    np.arange((1 << 20) - 64, (1 << 20) + 64), np.arange((1 << 16) - 64, (1 << 16) + 64),
                                    np.arange(3 * (1 << 16) - 64, 3 * (1 << 16) + 64)]))
    sample = st.reshape(n, 12)[idx].copy().reshape(-1, 4)
    exp = pr.orc.permute_batch(sample).reshape(-1, 12)
    assert np.array_equal(got[idx], exp)
    assert len(np.unique(got[idx], axis=0)) == len(idx)


def test_multi_chunk_soa_and_hash_host_path(S, pairs):
    """SoA layout and absorb/squeeze through the staged host path with more than one chunk
    (strided 2-D copies per chunk)."""
    pr = pairs("bn254_t3")
    n = (1 << 20) + 777
    st = pr.orc.synth(n * 3, seed=0xF3)
    got = unsoa(pr.ctx.permute_batch(soa(st, n, 3), layout=S.LAYOUT_SOA), n, 3).reshape(n, 12)
    idx = np.unique(np.concatenate([np.arange(0, 512), np.arange(n - 512, n), np.arange((1 << 20) - 32, (1 << 20) + 32),
                                    np.arange((1 << 16) - 32, (1 << 16) + 32), np.arange(5 * (1 << 16) - 32, 5 * (1 << 16) + 32)]))
    exp = pr.orc.permute_batch(st.reshape(n, 12)[idx].copy().reshape(-1, 4)).reshape(-1, 12)
    assert np.array_equal(got[idx], exp)
    x = st[: n * 2]
    h = pr.ctx.absorb_squeeze_batch(x, n, 2, 1).reshape(n, 4)
    hs = unsoa(pr.ctx.absorb_squeeze_batch(soa(x, n, 2), n, 2, 1, layout=S.LAYOUT_SOA), n, 1).reshape(n, 4)
    exp_h = pr.orc.absorb_squeeze_batch(x.reshape(n, 8)[idx].copy().reshape(-1, 4), len(idx), 2, 1).reshape(-1, 4)
    assert np.array_equal(h[idx], exp_h) and np.array_equal(hs[idx], exp_h)
