"""GPU parity for the single-sponge drop-in and the lock-step batch: the reference's own tests
(src/poseidon/mod.rs:376-399, src/poseidon/tests.rs:26-117) re-stated against the CUDA library,
plus state-machine fuzzing against the oracle (the reference has none)."""
import random

import numpy as np
import pytest

from conftest import field_id, field_name, load_golden, oracle_cfg, to_ints, to_limbs

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def S():
    import sponge_b200
    return sponge_b200


def gpu_cfg(S, ref, name):
    c = oracle_cfg(ref, name)
    return c, S.PoseidonConfig.new(c.full_rounds, c.partial_rounds, c.alpha, c.mds, c.ark, c.rate, c.capacity,
                                   field_name(ref, c))


@pytest.fixture(scope="module")
def fixture_pair(S, ref):
    o, g = gpu_cfg(S, ref, "bls_fixture")
    yield o, g
    g.context().close()


@pytest.fixture(scope="module")
def default_pair(S, ref):
    o, g = gpu_cfg(S, ref, "bls_default_r2")
    yield o, g
    g.context().close()


def test_poseidon_sponge_consistency(S):
    """reference src/poseidon/mod.rs:376-399, with parameters from the library's own
    get_default_poseidon_parameters (Grain LFSR in C++)."""
    kat = load_golden("reference_kats.json")["sponge_consistency"]
    param = S.get_default_poseidon_parameters("bls12_381", 2, False)
    sponge = S.PoseidonSponge.new(param)
    sponge.absorb([0, 1, 2])
    res = sponge.squeeze_native_field_elements(3)
    assert [str(v) for v in res] == kat["squeeze_native"]
    param.context().close()


def test_quirk_q1_squeeze_straddle(S, ref, default_pair):
    """mod.rs:175: remaining == rate is tested before the decrement."""
    g = load_golden("derived_vectors.json")["bls_default_q1"]
    _, cfg = default_pair
    s = S.PoseidonSponge(cfg)
    s.absorb([7])
    first = s.squeeze_native_field_elements(1)
    second = s.squeeze_native_field_elements(2)
    assert [str(v) for v in first] == g["first"] and [str(v) for v in second] == g["second"]
    assert second[1] == first[0]
    assert s.mode == S.DuplexSpongeMode.Squeezing(1)


def test_single_field_element_and_lists_differ(S, fixture_pair):
    """tests.rs:26-43 (single_field_element, list_with_constant_size_element)."""
    _, cfg = fixture_pair
    rng = random.Random(5)
    p = cfg.modulus
    e1 = rng.randrange(p)
    a, b = S.PoseidonSponge(cfg), S.PoseidonSponge(cfg)
    a.absorb(e1)
    b.absorb((e1 + 1) % p)
    assert a.squeeze_native_field_elements(3) != b.squeeze_native_field_elements(3)
    lst1 = [rng.randrange(p) for _ in range(1024 * 8)]
    lst2 = list(lst1)
    lst2[3] = (lst2[3] + 1) % p
    a, b = S.PoseidonSponge(cfg), S.PoseidonSponge(cfg)
    a.absorb(lst1)
    b.absorb(lst2)
    assert a.squeeze_native_field_elements(3) != b.squeeze_native_field_elements(3)


def test_long_absorb_matches_oracle(S, ref, fixture_pair):
    ocfg, cfg = fixture_pair
    rng = random.Random(6)
    lst = [rng.randrange(cfg.modulus) for _ in range(8192)]
    s = S.PoseidonSponge(cfg)
    s.absorb(lst)
    o = ref.PoseidonSponge(ocfg)
    o.absorb(lst)
    assert s.squeeze_native_field_elements(3) == o.squeeze_native_field_elements(3)


def test_squeeze_cast_native(S, fixture_pair):
    """tests.rs:71-85."""
    _, cfg = fixture_pair
    s1 = S.PoseidonSponge(cfg)
    s1.absorb(random.Random(7).randrange(cfg.modulus))
    s2 = s1.clone()
    assert s1.squeeze_native_field_elements(5) == s2.squeeze_field_elements(5, "bls12_381")


def test_split_absorb_equals_one_absorb(S, fixture_pair):
    """tests.rs:87-99 (absorb! == sequential absorb) and quirk Q2 (lazy permute)."""
    _, cfg = fixture_pair
    s1, s2 = S.PoseidonSponge(cfg), S.PoseidonSponge(cfg)
    s1.absorb([1, 2, 3, 4, 5, 6])
    s1.absorb(114514)
    s2.absorb([1, 2, 3, 4, 5, 6, 114514])
    assert s1.squeeze_native_field_elements(3) == s2.squeeze_native_field_elements(3)
    s3 = S.PoseidonSponge(cfg)
    s3.absorb([1, 2])
    assert s3.mode == S.DuplexSpongeMode.Absorbing(2)     # rate filled, no permute yet
    assert s3.state[1:] == [1, 2] and s3.state[0] == 0


def test_empty_absorb_is_noop_and_zero_squeeze_permutes(S, ref, fixture_pair):
    ocfg, cfg = fixture_pair
    s = S.PoseidonSponge(cfg)
    s.absorb([])
    assert s.mode == S.DuplexSpongeMode.Absorbing(0) and s.state == [0, 0, 0]
    assert s.squeeze_native_field_elements(0) == []
    assert s.mode == S.DuplexSpongeMode.Squeezing(0)       # mod.rs:320-342: permutes and flips
    assert s.state == ref.permute(ocfg, [0, 0, 0])


@pytest.mark.parametrize("name", ["bls_fixture", "bn254_t3", "bn254_t5", "bn254_t4_cap2", "bls_default_r8"])
def test_state_machine_fuzz(S, ref, name):
    """Random call sequences: absorb(k) / squeeze_native(k) / squeeze_bytes / clone / state export."""
    ocfg, cfg = gpu_cfg(S, ref, name)
    rng = random.Random(hash(name) & 0xffff)
    p = cfg.modulus
    try:
        for trial in range(6):
            s, o = S.PoseidonSponge(cfg), ref.PoseidonSponge(ocfg)
            for step in range(14):
                op = rng.choice(["absorb", "absorb", "squeeze", "bytes", "clone", "state"])
                if op == "absorb":
                    xs = [rng.randrange(p) for _ in range(rng.choice([0, 1, 1, 2, 3, cfg.rate, cfg.rate + 1, 11]))]
                    s.absorb(xs)
                    o.absorb(xs)
                elif op == "squeeze":
                    k = rng.choice([0, 1, 1, 2, cfg.rate, cfg.rate + 1, 2 * cfg.rate, 7])
                    assert s.squeeze_native_field_elements(k) == o.squeeze_native_field_elements(k)
                elif op == "bytes":
                    k = rng.choice([0, 1, 31, 32, 62, 63, 100])
                    assert s.squeeze_bytes(k) == o.squeeze_bytes(k)
                elif op == "clone":
                    s, o = s.clone(), o.clone()
                else:
                    st = s.into_state()
                    assert st.state == o.state
                    assert (st.mode.kind == "Absorbing") == (o.mode == ref.ABSORBING) and st.mode.index == o.index
                    s = S.PoseidonSponge.from_state(st, cfg)
            assert s.squeeze_native_field_elements(2) == o.squeeze_native_field_elements(2)
    finally:
        cfg.context().close()


def test_bits_nonnative_bytes_fork_golden(S, ref, fixture_pair):
    g = load_golden("derived_vectors.json")
    ocfg, cfg = fixture_pair
    s = S.PoseidonSponge(cfg)
    s.absorb([1, 2, 3])
    bits = s.squeeze_bits(300)
    assert "".join("1" if b else "0" for b in bits) == g["bls_fixture_absorb_1_2_3_squeeze_bits_300"]
    s = S.PoseidonSponge(cfg)
    s.absorb([1, 2, 3])
    assert [str(v) for v in s.squeeze_field_elements(3, "bn254")] == g["bls_fixture_absorb_1_2_3_squeeze_bn254_3"]
    s = S.PoseidonSponge(cfg)
    s.absorb(bytes(range(100)))
    assert [str(v) for v in s.squeeze_native_field_elements(2)] == g["bls_fixture_absorb_bytes_0_99_squeeze_2"]
    s = S.PoseidonSponge(cfg)
    s.absorb(5)
    f = s.fork(b"domain-A")
    assert [str(v) for v in f.squeeze_native_field_elements(2)] == g["bls_fixture_absorb_5_fork_domainA_squeeze_2"]
    # the parent is untouched by fork
    o = ref.PoseidonSponge(ocfg)
    o.absorb([5])
    assert s.squeeze_native_field_elements(1) == o.squeeze_native_field_elements(1)


def test_sizes_quirk_q3(S, fixture_pair):
    """lib.rs:45-52: Truncated(n) is validated but has no effect on the output width."""
    _, cfg = fixture_pair
    a, b = S.PoseidonSponge(cfg), S.PoseidonSponge(cfg)
    a.absorb([9])
    b.absorb([9])
    full = a.squeeze_field_elements_with_sizes([S.FieldElementSize.Full()] * 2, "bn254")
    trunc = b.squeeze_field_elements_with_sizes([S.FieldElementSize.Truncated(10)] * 2, "bn254")
    assert full == trunc
    with pytest.raises(ValueError):
        S.FieldElementSize.Truncated(300).num_bits("bn254")


def test_sponge_batch_lockstep(S, ref):
    """psd_batch: n sponges, several absorb/squeeze calls, against n independent oracle sponges."""
    from oracle.c_oracle import COracle, COracleSponge
    ocfg, cfg = gpu_cfg(S, ref, "bn254_t3")
    ctx = cfg.context()
    try:
        orc = COracle(ocfg, field_id(ref, ocfg))
        n = 257
        b = S.SpongeBatch(ctx, n)
        os_ = [COracleSponge(orc) for _ in range(n)]
        seq = [("a", 1), ("a", 3), ("s", 1), ("s", 2), ("a", 2), ("a", 5), ("s", 4), ("s", 0), ("a", 1), ("s", 3)]
        for i, (op, k) in enumerate(seq):
            if op == "a":
                x = orc.synth(n * k, seed=0x50 + i)
                b.absorb(x, k)
                for j, o in enumerate(os_):
                    o.absorb(x[j * k:(j + 1) * k])
            else:
                got = b.squeeze_native(k)
                exp = np.stack([o.squeeze(k) for o in os_])    # squeeze(0) still permutes (mod.rs:320-342)
                assert np.array_equal(got, exp.reshape(n, k, 4))
        c = b.clone()
        assert np.array_equal(b.squeeze_bytes(40), c.squeeze_bytes(40))
        st, mode = b.get_state(layout=S.LAYOUT_AOS, repr=S.REPR_MONTGOMERY)
        exp_st = np.stack([o.state()[0] for o in os_])
        assert np.array_equal(st, exp_st)
        assert mode.kind == "Squeezing" and mode.index == os_[0].state()[2]
        # import into a fresh batch and continue
        b2 = S.SpongeBatch(ctx, n)
        b2.set_state(st, mode)
        assert np.array_equal(b2.squeeze_native(2), b.squeeze_native(2))
    finally:
        ctx.close()


def test_cpp_wrapper_reference_kat(S):
    """The C++ mirror (include/poseidon_b200.hpp) reproduces src/poseidon/mod.rs:376-399 on the GPU."""
    import subprocess
    from test_host_logic import _build_cpp_wrapper_test
    exe = _build_cpp_wrapper_test(S)
    kat = load_golden("reference_kats.json")["sponge_consistency"]["squeeze_native"]
    r = subprocess.run([exe] + ["%064x" % int(v) for v in kat], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr


def test_typed_absorb_and_macros(S, ref, fixture_pair):
    """tests.rs:87-102 (absorb! == sequential absorb) and typed inputs through sponge.absorb:
    ints, signed ints, byte slices, Option -- against the oracle fed with the same field elements."""
    from sponge_b200 import absorb as A
    ocfg, cfg = fixture_pair
    s1, s2 = S.PoseidonSponge(cfg), S.PoseidonSponge(cfg)
    v = [A.I32(x) for x in (1, 2, 3, 4, 5, 6)]
    s1.absorb(v)
    s1.absorb(114514)
    A.absorb_all(s2, v, A.Fe(114514))
    assert s1.squeeze_native_field_elements(3) == s2.squeeze_native_field_elements(3)
    s3, o = S.PoseidonSponge(cfg), ref.PoseidonSponge(ocfg)
    for item in ([A.I16(-7), A.U64(9)], bytes(range(40)), A.Some(A.U8(3)), A.NONE, True):
        s3.absorb(item)
        o.absorb(A.to_sponge_field_elements(item, "bls12_381"))
    assert s3.squeeze_native_field_elements(2) == o.squeeze_native_field_elements(2)
