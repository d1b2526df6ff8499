"""GPU parity for kernel variants 6 and 7 (integer S-box + the whole linear layer on the tensor cores: byte-digit tables,
u8 x u8 -> s32, one 32-bit Montgomery row per output value; 6 = mma.sync with register fragments, 7 = tcgen05.mma with
operands in shared memory and accumulators in TMEM): bit-exact against the CPU oracle through the C ABI, for every parameter shape the variant is built for, batch sizes that leave the last warp partly filled
(the kernel keeps padding threads alive for the warp-collective instructions), every layout / representation,
the sponge call sequences and the Merkle driver."""
import numpy as np
import pytest

from conftest import field_id, field_name, oracle_cfg, to_limbs

pytestmark = pytest.mark.gpu

NAMES = ["bn254_t3", "bn254_t2", "bn254_t5", "bn254_t4_cap2", "bn254_full_only", "bn254_rp1", "bn254_rp2_alpha17",
         "bls_fixture", "bls_default_r2", "bls_t3_alpha5"]


def make(ref, name):
    import sponge_b200 as S
    from oracle.c_oracle import COracle
    oc = oracle_cfg(ref, name)
    cfg = S.PoseidonConfig(oc.full_rounds, oc.partial_rounds, oc.alpha, oc.mds, oc.ark, oc.rate, oc.capacity,
                           field_name(ref, oc))
    return S, oc, COracle(oc, field_id(ref, oc)), cfg.context()


def has_variant7(oc):
    """Variant 7 (tcgen05 / TMEM form of the same contraction) is built for the same widths and exponents as variant 6."""
    return oc.rate + oc.capacity <= 5


@pytest.mark.parametrize("variant", [6, 7])
@pytest.mark.parametrize("name", NAMES)
def test_tc_permute_matches_oracle(ref, name, variant):
    S, oc, orc, ctx = make(ref, name)
    try:
        if variant == 7 and not has_variant7(oc):
            pytest.skip("variant 7 is not built for this width")
        ctx.set_variant(variant)
        assert ctx.active_variant() == variant, ctx.describe()
        t = oc.rate + oc.capacity
        for n in (1, 31, 33, 127, 129, 4097, 20000, 148 * 512 + 77):
            st = orc.synth(n * t, seed=0x7C0 + n)
            if n >= 33:
                st[:t] = 0
                st[t:2 * t] = orc.to_mont(to_limbs([oc.p - 1] * t))
                st[2 * t:3 * t] = to_limbs([oc.p - 1] * t)          # largest Montgomery images
            exp = orc.permute_batch(st.copy())
            assert np.array_equal(ctx.permute_batch(st.copy()), exp), (name, n)
        n = 3000
        st = orc.synth(n * t, seed=0x7C1)
        exp = orc.permute_batch(st.copy())
        canon = orc.from_mont(st.copy())
        assert np.array_equal(ctx.permute_batch(canon, repr=S.REPR_CANONICAL), orc.from_mont(exp.copy()))
        soa = np.ascontiguousarray(st.reshape(n, t, 2, 2).transpose(1, 2, 0, 3))
        got = ctx.permute_batch(soa, layout=S.LAYOUT_SOA)
        assert np.array_equal(np.ascontiguousarray(got.reshape(t, 2, n, 2).transpose(2, 0, 1, 3)).reshape(-1, 4), exp)
        # the other variants agree
        for v in (5, 3, 0):
            ctx.set_variant(v)
            assert np.array_equal(ctx.permute_batch(st.copy()), exp), (name, v, ctx.describe())
    finally:
        ctx.close()


@pytest.mark.parametrize("variant", [6, 7])
@pytest.mark.parametrize("name", ["bn254_t3", "bn254_t5", "bls_fixture"])
def test_tc_sponge_calls_and_merkle(ref, name, variant):
    S, oc, orc, ctx = make(ref, name)
    try:
        if variant == 7 and not has_variant7(oc):
            pytest.skip("variant 7 is not built for this width")
        ctx.set_variant(variant)
        assert ctx.active_variant() == variant
        for n, n_absorb, n_squeeze in ((700, 5, 3), (45, 1, 1), (1000, 0, 2), (513, 9, 4)):
            x = orc.synth(max(n * n_absorb, 1), seed=0x7C2 + n_absorb)[: n * n_absorb]
            assert np.array_equal(ctx.absorb_squeeze_batch(x, n, n_absorb, n_squeeze),
                                  orc.absorb_squeeze_batch(x, n, n_absorb, n_squeeze)), (n, n_absorb, n_squeeze)
        x = orc.synth(300 * 7, seed=0x7C3)
        assert np.array_equal(ctx.absorb_squeeze_bytes_batch(x, 300, 7, 50), orc.absorb_squeeze_bytes_batch(x, 300, 7, 50))
        if oc.rate >= 2:
            leaves = orc.synth(1 << 12, seed=0x7C4)
            assert np.array_equal(ctx.merkle_root(leaves), orc.merkle_root(leaves))
        # lock-step batch with state kept on the device between calls
        b = S.SpongeBatch(ctx, 77)
        xs = orc.synth(77 * 6, seed=0x7C5)
        b.absorb(np.ascontiguousarray(xs.reshape(77, 6, 4)[:, :2]), 2)
        b.absorb(np.ascontiguousarray(xs.reshape(77, 6, 4)[:, 2:]), 4)
        assert np.array_equal(b.squeeze_native(3), orc.absorb_squeeze_batch(xs, 77, 6, 3))
        b.free()
        # the single-sponge drop-in (a batch of one: 31 padding threads in its warp)
        s = S.PoseidonSponge(ctx.cfg)
        s.absorb([1, 2, 3])
        from oracle import poseidon_ref as R
        assert s.squeeze_native_field_elements(2) == R.hash_absorb_squeeze(oc, [1, 2, 3], 2)
    finally:
        ctx.close()
