"""Many live parameter sets on one GPU.  The reference lets a program hold any number of PoseidonConfig
values and sponges at once (a sponge owns a clone of its parameters, poseidon/mod.rs:44-60, 219-230); here
constants live either in the constant bank (one parameter set per kernel unit, more or less) or in the
kernels' shared-memory form.  Checked through the C ABI against the oracle:
  * three and more live contexts of the same field and width, different parameters: every one is created,
    says where its constants live and which kernel it runs, and is bit-exact;
  * identical parameter sets share one image;
  * the second BN254 t=3 parameter set runs the same algorithm as the first, at nearly the same rate;
  * the whole parity suite again with every context forced into the shared-memory form."""
import os
import subprocess
import sys
import time

import numpy as np
import pytest

from conftest import field_id, field_name

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def make_cfg(S, ref, p, rate, rf, rp, alpha, skip):
    ark, mds = ref.find_poseidon_ark_and_mds(p, 254 if p == ref.BN254_FR else 255, rate, rf, rp, skip)
    ocfg = ref.PoseidonConfig(rf, rp, alpha, mds, ark, rate, 1, p)
    cfg = S.PoseidonConfig(rf, rp, alpha, mds, ark, rate, 1, field_name(ref, ocfg))
    return ocfg, cfg


@pytest.mark.parametrize("field,rate,rf,rp,alpha", [("bn254", 2, 8, 57, 5), ("bn254", 4, 8, 60, 5), ("bls", 2, 8, 29, 17),
                                                    ("bn254", 3, 8, 56, 5), ("bls", 7, 8, 20, 5)])
def test_many_live_parameter_sets(ref, field, rate, rf, rp, alpha):
    import sponge_b200 as S
    from oracle.c_oracle import COracle
    p = ref.BN254_FR if field == "bn254" else ref.BLS12_381_FR
    t = rate + 1
    live = []
    try:
        for skip in range(4):   # four different MDS matrices -> four different parameter sets, all alive
            ocfg, cfg = make_cfg(S, ref, p, rate, rf, rp, alpha, skip)
            live.append((ocfg, S.PoseidonContext(cfg, 0)))
        variants = [c.describe()["variant"] for _, c in live]
        assert len(set(variants)) == 1, f"parameter sets of one shape run different algorithms by default: {variants}"
        for _, c in live:
            c.set_variant(5)   # the forms whose constants compete for the constant bank
        homes = [c.describe()["constants_home"] for _, c in live]
        assert homes[0] == "bank" or (field == "bn254" and t == 5), homes   # t=5 variant 5 exists in the shared form only
        if t in (3, 5):
            assert "shared" in homes, homes                                  # the bank cannot hold four of these
        variants = [c.describe()["variant"] for _, c in live]
        assert len(set(variants)) == 1, f"parameter sets of one shape run different algorithms: {variants}"
        for k, (ocfg, ctx) in enumerate(live):
            orc = COracle(ocfg, field_id(ref, ocfg))
            st = orc.synth(1500 * t, seed=0xD0 + k)
            assert np.array_equal(ctx.permute_batch(st.copy()), orc.permute_batch(st.copy())), (k, ctx.describe())
            x = orc.synth(700 * 5, seed=0xE0 + k)
            assert np.array_equal(ctx.absorb_squeeze_batch(x, 700, 5, 2), orc.absorb_squeeze_batch(x, 700, 5, 2))
        # every variant request on a shared-memory context still gives the same permutation
        ocfg, ctx = live[-1]
        orc = COracle(ocfg, field_id(ref, ocfg))
        st = orc.synth(600 * t, seed=0xF1)
        exp = orc.permute_batch(st.copy())
        for v in (0, 1, 2, 3, 5, 6, 7):
            ctx.set_variant(v)
            assert np.array_equal(ctx.permute_batch(st.copy()), exp), (v, ctx.describe())
    finally:
        for _, c in live:
            c.close()


def test_identical_parameter_sets_share_one_image(ref):
    import sponge_b200 as S
    ocfg, cfg = make_cfg(S, ref, ref.BN254_FR, 2, 8, 57, 5, 0)
    a = S.PoseidonContext(cfg, 0)
    b = S.PoseidonContext(cfg, 0)
    c = S.PoseidonContext(cfg, 0)
    try:
        assert a.describe()["image_shared_by"] == 3          # whatever the default kernel reads is shared too
        for x in (a, b, c):
            x.set_variant(5)
        da, dc = a.describe(), c.describe()
        assert dc["image_shared_by"] == 3 and da["constants_home"] == dc["constants_home"] == "bank"
        assert da["kernel_name"] == dc["kernel_name"]
        b.close()
        assert a.describe()["image_shared_by"] == 2
    finally:
        for x in (a, b, c):
            x.close()
    # the slot is free again: a new parameter set gets the bank
    ocfg2, cfg2 = make_cfg(S, ref, ref.BN254_FR, 2, 8, 57, 5, 1)
    d = S.PoseidonContext(cfg2, 0)
    try:
        d.set_variant(5)
        assert d.describe()["constants_home"] == "bank"
    finally:
        d.close()


def test_second_parameter_set_is_not_a_performance_cliff(ref):
    """Two live BN254 t=3 parameter sets: both report what they run and get the same default kernel; with variant 5
    requested (the form whose tables compete for the constant bank) both run variant 5, and the one in the
    shared-memory form is within 10 % of the one in the constant bank on 2^20 permutations."""
    import torch
    import sponge_b200 as S
    _, cfg1 = make_cfg(S, ref, ref.BN254_FR, 2, 8, 57, 5, 0)
    _, cfg2 = make_cfg(S, ref, ref.BN254_FR, 2, 8, 57, 5, 1)
    c1, c2 = S.PoseidonContext(cfg1, 0), S.PoseidonContext(cfg2, 0)
    try:
        assert c1.describe()["variant"] == c2.describe()["variant"]   # the default is the same for both
        c1.set_variant(5)
        c2.set_variant(5)
        d1, d2 = c1.describe(), c2.describe()
        assert d1["variant"] == d2["variant"] == 5
        assert (d1["constants_home"], d2["constants_home"]) == ("bank", "shared")
        n = 1 << 20
        buf = torch.empty((n * 3, 4), dtype=torch.int64, device="cuda")
        s = torch.cuda.current_stream().cuda_stream
        c1.synth_fill_dev(buf.data_ptr(), n * 3, 7, stream=s)
        rates = []
        for ctx in (c1, c2):
            for _ in range(2):
                ctx.permute_batch_dev(buf.data_ptr(), n, stream=s)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(5):
                ctx.permute_batch_dev(buf.data_ptr(), n, stream=s)
            e1.record()
            torch.cuda.synchronize()
            rates.append(5 * n / (e0.elapsed_time(e1) * 1e-3))
        print(f"bank {rates[0]:.4g} perm/s, shared {rates[1]:.4g} perm/s, ratio {rates[1] / rates[0]:.3f}")
        assert rates[1] > 0.90 * rates[0], rates
    finally:
        c1.close()
        c2.close()


def test_parity_suite_with_every_context_in_the_shared_memory_form():
    """PSD_FORCE_SMEM=1 makes psd_ctx_create skip the constant bank; the permutation / absorb-squeeze /
    variant-5 parity tests then exercise the shared-memory kernels of every field and width."""
    env = dict(os.environ, PSD_FORCE_SMEM="1")
    sel = ["tests/test_gpu_parity.py::test_permute_batch_matches_oracle",
           "tests/test_gpu_parity.py::test_permute_dense_variant_equals_optimised",
           "tests/test_gpu_parity.py::test_absorb_squeeze_bytes_batch",
           "tests/test_gpu_fp52.py"]
    t0 = time.time()
    r = subprocess.run([sys.executable, "-m", "pytest", "-x", "-q", "-m", "gpu", "-p", "no:cacheprovider"] + sel,
                       cwd=ROOT, env=env, capture_output=True, text=True, timeout=1500)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-2000:]
    print(f"forced shared-memory parity run: {r.stdout.strip().splitlines()[-1]} ({time.time() - t0:.0f} s)")
