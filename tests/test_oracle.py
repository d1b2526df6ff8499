"""The oracle is pinned before it is trusted: every known-answer value the reference's own tests
hold for this path (33 values, tests/golden/reference_kats.json), then the C restatement against the
Python one on random inputs.  CPU only."""
import random

import numpy as np
import pytest

from conftest import field_id, load_golden, oracle_cfg, to_ints, to_limbs


def test_grain_lfsr_kat(ref):
    """reference src/poseidon/grain_lfsr.rs:197-213"""
    k = load_golden("reference_kats.json")["grain_lfsr"]
    lfsr = ref.GrainLFSR(*k["args"])
    a = lfsr.field_elements_rejection_sampling(1, ref.BLS12_381_FR)[0]
    b = lfsr.field_elements_rejection_sampling(1, ref.BLS12_381_FR)[0]
    assert [str(a), str(b)] == k["rejection_samples"]


def test_default_parameters_kats(ref):
    """reference src/poseidon/traits.rs:163-358: 14 parameter sets x (ark[0][0], mds[0][0])"""
    entries = load_golden("reference_kats.json")["default_parameters"]["entries"]
    assert len(entries) == 14
    for e in entries:
        c = ref.get_default_poseidon_parameters(e["rate"], e["optimized_for_weights"])
        assert str(c.ark[0][0]) == e["ark00"] and str(c.mds[0][0]) == e["mds00"]
        assert c.capacity == 1 and len(c.ark) == c.full_rounds + c.partial_rounds
    assert ref.get_default_poseidon_parameters(9, False) is None


def test_sponge_consistency_kat(ref):
    """reference src/poseidon/mod.rs:376-399"""
    k = load_golden("reference_kats.json")["sponge_consistency"]
    cfg = ref.get_default_poseidon_parameters(2, False)
    s = ref.PoseidonSponge(cfg)
    s.absorb([0, 1, 2])
    assert [str(v) for v in s.squeeze_native_field_elements(3)] == k["squeeze_native"]
    assert s.n_permutes == 3


def test_fixture_params_shape(fixture_cfg, ref):
    """poseidon_parameters_for_test (tests.rs:120-820): alpha 17, 8 + 29 rounds, all values < p."""
    c = fixture_cfg
    assert (c.alpha, c.full_rounds, c.partial_rounds, c.rate, c.capacity) == (17, 8, 29, 2, 1)
    assert all(0 <= v < c.p for row in c.ark + c.mds for v in row)
    # it is not LFSR-derivable: a literal table
    ark, _ = ref.find_poseidon_ark_and_mds(c.p, 255, 2, 8, 29, 0)
    assert ark[0][0] != c.ark[0][0]


def test_external_circomlib_vector(ref):
    c = oracle_cfg(ref, "bn254_t3")
    assert ref.permute(c, [0, 1, 2])[0] == 0x115cc0f5e7d690413df64c6b9662e9cf2a3617f2743245519e19607a4417189a


def test_derived_golden_regression(ref, fixture_cfg):
    g = load_golden("derived_vectors.json")
    bn3, bn5 = oracle_cfg(ref, "bn254_t3"), oracle_cfg(ref, "bn254_t5")
    assert "0x%064x" % ref.hash_absorb_squeeze(bn3, [1, 2], 1)[0] == g["bn254_t3_absorb_1_2_squeeze_1"]
    assert "0x%064x" % ref.hash_absorb_squeeze(bn5, [1, 2, 3, 4], 1)[0] == g["bn254_t5_absorb_1_2_3_4_squeeze_1"]
    assert "0x%064x" % ref.hash_absorb_squeeze(fixture_cfg, [1, 2], 1)[0] == g["bls_fixture_absorb_1_2_squeeze_1"]
    assert "0x%064x" % ref.merkle_root(bn3, list(range(8))) == g["bn254_t3_merkle_leaves_0_7"]


def test_state_machine_quirks(ref):
    cfg = ref.get_default_poseidon_parameters(2, False)
    # Q1: squeeze(1) then squeeze(2) -> one permutation in total, second[1] == first[0]
    s = ref.PoseidonSponge(cfg)
    s.absorb([7])
    first = s.squeeze_native_field_elements(1)
    second = s.squeeze_native_field_elements(2)
    assert s.n_permutes == 1 and second[1] == first[0]
    s2 = ref.PoseidonSponge(cfg)
    s2.absorb([7])
    assert s2.squeeze_native_field_elements(3)[:1] == first and s2.n_permutes == 2
    # Q2: lazy absorb + split absorb == single absorb
    a, b = ref.PoseidonSponge(cfg), ref.PoseidonSponge(cfg)
    a.absorb([1, 2])
    assert a.n_permutes == 0 and (a.mode, a.index) == (ref.ABSORBING, 2)
    a.absorb([3])
    b.absorb([1, 2, 3])
    assert a.squeeze_native_field_elements(2) == b.squeeze_native_field_elements(2)
    # empty absorb is a no-op; squeeze(0) still permutes
    e = ref.PoseidonSponge(cfg)
    e.absorb([])
    assert e.n_permutes == 0
    assert e.squeeze_native_field_elements(0) == [] and e.n_permutes == 1 and e.mode == ref.SQUEEZING


def test_montgomery_properties(ref):
    for p in (ref.BN254_FR, ref.BLS12_381_FR):
        rng = random.Random(p & 0xffff)
        for _ in range(50):
            x = rng.randrange(p)
            assert ref.from_mont(ref.to_mont(x, p), p) == x
        assert ref.to_mont(1, p) == (1 << 256) % p


CFGS = ["bn254_t3", "bn254_t5", "bls_default_r2", "bls_fixture", "bls_default_r3", "bls_default_r2_w",
        "bn254_t2", "bn254_t4_cap2", "bn254_odd_rf", "bls_default_r8"]


@pytest.mark.parametrize("name", CFGS)
def test_c_oracle_matches_python_oracle(ref, name):
    from oracle.c_oracle import COracle, COracleSponge
    cfg = oracle_cfg(ref, name)
    orc = COracle(cfg, field_id(ref, cfg))
    rng = random.Random(11)
    p, t = cfg.p, cfg.rate + cfg.capacity
    # permutation
    sts = [[rng.randrange(p) for _ in range(t)] for _ in range(6)] + [[0] * t, [p - 1] * t]
    a = orc.to_mont(to_limbs(v for s in sts for v in s))
    got = to_ints(orc.from_mont(orc.permute_batch(a)))
    assert got == [v for s in sts for v in ref.permute(cfg, s)]
    # one-shot hashes
    for n_abs, n_sq in [(2, 1), (0, 1), (1, 0), (cfg.rate, cfg.rate), (cfg.rate + 1, 2 * cfg.rate + 1), (9, 4)]:
        ins = [[rng.randrange(p) for _ in range(n_abs)] for _ in range(3)]
        x = orc.to_mont(to_limbs(v for s in ins for v in s)) if n_abs else np.zeros((0, 4), dtype=np.uint64)
        out = orc.absorb_squeeze_batch(x, 3, n_abs, n_sq)
        assert to_ints(orc.from_mont(out)) == [v for s in ins for v in ref.hash_absorb_squeeze(cfg, s, n_sq)]
    # call-by-call state machine incl. Q1
    o, s = ref.PoseidonSponge(cfg), COracleSponge(orc)
    for step in range(25):
        if rng.random() < 0.5:
            xs = [rng.randrange(p) for _ in range(rng.choice([0, 1, 2, cfg.rate, cfg.rate + 1, 5]))]
            o.absorb(xs)
            s.absorb(orc.to_mont(to_limbs(xs)) if xs else np.zeros((0, 4), dtype=np.uint64))
        else:
            k = rng.choice([0, 1, 2, cfg.rate, cfg.rate + 1, 2 * cfg.rate])
            assert to_ints(orc.from_mont(s.squeeze(k).copy())) == o.squeeze_native_field_elements(k)
        st, mode, idx = s.state()
        assert to_ints(orc.from_mont(st.copy())) == o.state and (mode, idx) == (o.mode, o.index)


def test_c_oracle_bytes_merkle_synth(ref, fixture_cfg):
    from oracle.c_oracle import COracle
    g = load_golden("derived_vectors.json")
    orc = COracle(fixture_cfg, 1)
    x = orc.to_mont(to_limbs(range(1024)))
    assert bytes(orc.absorb_squeeze_bytes_batch(x, 1, 1024, 32)[0]).hex() == g["bls_fixture_absorb_0_1023_squeeze_bytes_32"]
    bn3 = oracle_cfg(ref, "bn254_t3")
    ob = COracle(bn3, 0)
    root = ob.from_mont(ob.merkle_root(ob.to_mont(to_limbs(range(1, 1025)))))
    assert "0x%064x" % to_ints(root)[0] == g["bn254_t3_merkle_leaves_1_1024"]
    # Montgomery limbs of a known value
    h = ob.absorb_squeeze_batch(ob.to_mont(to_limbs([1, 2])), 1, 2, 1)
    assert ["0x%016x" % v for v in h.reshape(-1)] == g["bn254_t3_absorb_1_2_squeeze_1_montgomery_limbs64"]
    # synthetic generator: C == Python, canonical and uniform-looking
    s = ob.synth(64, 0xC2, 1000, montgomery=False)
    assert to_ints(s) == [ref.synth_element(0xC2, 1000 + k, ref.BN254_FR) for k in range(64)]
    assert np.array_equal(ob.synth(64, 0xC2, 1000, montgomery=True), ob.to_mont(s.copy()))
    # OpenMP result does not depend on the thread count
    st = ob.synth(3 * 512, 9)
    assert np.array_equal(ob.permute_batch(st.copy(), threads=1), ob.permute_batch(st.copy(), threads=4))
