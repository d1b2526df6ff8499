"""GPU parity for kernel variant 5 (integer S-box + linear layer on the FP64 pipe in 52-bit limbs):
bit-exact against the CPU oracle and against the integer variants, through the C ABI.  Each context is
created on its own so that it is the one holding the variant's constant pool; the test fails if the
library silently ran another variant."""
import numpy as np
import pytest

from conftest import field_id, field_name, oracle_cfg, to_limbs

pytestmark = pytest.mark.gpu

NAMES = ["bn254_t3", "bn254_t2", "bn254_full_only", "bn254_rp1", "bn254_rp2_alpha17", "bn254_t5"]


@pytest.fixture(scope="module")
def S():
    import sponge_b200
    return sponge_b200


def make(S, ref, name):
    from oracle.c_oracle import COracle
    oc = oracle_cfg(ref, name)
    orc = COracle(oc, field_id(ref, oc))
    cfg = S.PoseidonConfig(oc.full_rounds, oc.partial_rounds, oc.alpha, oc.mds, oc.ark, oc.rate, oc.capacity,
                           field_name(ref, oc))
    return oc, orc, cfg.context()


@pytest.mark.parametrize("name", NAMES)
def test_fp52_permute_matches_oracle(S, ref, name):
    oc, orc, ctx = make(S, ref, name)
    try:
        ctx.set_variant(5)
        assert ctx.active_variant() == 5, "the variant-5 pool was not given to this context"
        t = oc.rate + oc.capacity
        n = 5000
        st = orc.synth(n * t, seed=0xF52, montgomery=True)
        st[:t] = 0
        st[t:2 * t] = orc.to_mont(to_limbs([oc.p - 1] * t))
        st[2 * t:3 * t] = orc.to_mont(to_limbs([1] * t))
        exp = orc.permute_batch(st.copy())
        got = ctx.permute_batch(st.copy())
        assert np.array_equal(got, exp)
        # raw residues p-1 (largest Montgomery images) and 2^k - 1 patterns in every lane
        edge = to_limbs([oc.p - 1, (1 << 253) - 1, (1 << 208) - 1, (1 << 52) - 1, 1 << 52, (1 << 104) - 1] * t)[: 6 * t]
        assert np.array_equal(ctx.permute_batch(edge.copy()), orc.permute_batch(edge.copy()))
        # the integer variants agree
        ctx.set_variant(3)
        assert ctx.active_variant() != 5
        assert np.array_equal(ctx.permute_batch(st.copy()), exp)
    finally:
        ctx.close()


@pytest.mark.parametrize("n", [40000, 80000, 80001])
def test_fp52_large_blocks_match_oracle(S, ref, n):
    """Batches large enough for the 256- and 512-thread block launches (more than 256 / 512 sponges per
    SM), with a ragged last block, against the oracle element for element."""
    oc, orc, ctx = make(S, ref, "bn254_t3")
    try:
        ctx.set_variant(5)
        assert ctx.active_variant() == 5
        t = oc.rate + oc.capacity
        st = orc.synth(n * t, seed=0xB10C + n, montgomery=True)
        assert np.array_equal(ctx.permute_batch(st.copy()), orc.permute_batch(st.copy()))
    finally:
        ctx.close()


def test_fp52_sponge_calls(S, ref):
    """absorb/squeeze sequences and the Merkle driver run on the same kernel."""
    oc, orc, ctx = make(S, ref, "bn254_t3")
    try:
        ctx.set_variant(5)
        assert ctx.active_variant() == 5
        n = 4096
        inputs = orc.synth(n * 5, seed=0xF53, montgomery=True)
        got = ctx.absorb_squeeze_batch(inputs, n, n_absorb=5, n_squeeze=3)
        ctx.set_variant(2)
        exp = ctx.absorb_squeeze_batch(inputs, n, n_absorb=5, n_squeeze=3)
        assert np.array_equal(got, exp)
        leaves = orc.synth(1 << 12, seed=0xF54, montgomery=True)
        root2 = ctx.merkle_root(leaves)
        ctx.set_variant(5)
        assert np.array_equal(ctx.merkle_root(leaves), root2)
    finally:
        ctx.close()


def test_fp52_second_context_keeps_the_variant(S, ref):
    """The variant-5 constant pool holds one parameter set; the next one runs the same variant with its
    tables staged through shared memory instead of dropping to an integer variant."""
    import os
    oc, orc, ctx = make(S, ref, "bn254_t3")
    oc2, orc2, ctx2 = make(S, ref, "bn254_t2")
    try:
        ctx.set_variant(5)
        ctx2.set_variant(5)
        assert ctx.active_variant() == 5 and ctx2.active_variant() == 5
        if os.environ.get("PSD_FORCE_SMEM") != "1":
            assert ctx.describe()["constants_home"] == "bank" and ctx2.describe()["constants_home"] == "shared"
        st = orc2.synth(64 * 2, seed=3, montgomery=True)
        assert np.array_equal(ctx2.permute_batch(st.copy()), orc2.permute_batch(st.copy()))
    finally:
        ctx.close()
        ctx2.close()
