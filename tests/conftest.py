import json
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def load_golden(name):
    with open(os.path.join(GOLDEN, name)) as f:
        return json.load(f)


def to_limbs(values):
    vals = [int(v) for v in values]
    return np.frombuffer(b"".join(v.to_bytes(32, "little") for v in vals), dtype=np.uint64).reshape(len(vals), 4).copy()


def to_ints(arr):
    raw = np.ascontiguousarray(arr, dtype=np.uint64).tobytes()
    return [int.from_bytes(raw[i:i + 32], "little") for i in range(0, len(raw), 32)]


@pytest.fixture(scope="session")
def ref():
    from oracle import poseidon_ref
    return poseidon_ref


@pytest.fixture(scope="session")
def fixture_cfg(ref):
    """poseidon_parameters_for_test (reference src/poseidon/tests.rs:120-820) as an oracle config."""
    fx = load_golden("bls12_381_fixture_params.json")
    return ref.PoseidonConfig(fx["full_rounds"], fx["partial_rounds"], fx["alpha"],
                              [[int(v) for v in r] for r in fx["mds"]], [[int(v) for v in r] for r in fx["ark"]],
                              fx["rate"], fx["capacity"], ref.BLS12_381_FR)


_CFG_CACHE = {}


def oracle_cfg(ref, name, fixture=None):
    """Named parameter sets used across the tests (oracle-side objects)."""
    if name not in _CFG_CACHE:
        if name == "bn254_t3":       # BASELINE.json config 2
            c = ref.bn254_config(2, 8, 57)
        elif name == "bn254_t5":     # config 3
            c = ref.bn254_config(4, 8, 60)
        elif name == "bls_default_r2":   # the KAT-pinned default set, alpha 17
            c = ref.get_default_poseidon_parameters(2, False)
        elif name == "bls_fixture":  # configs 1 and 5
            fx = load_golden("bls12_381_fixture_params.json")
            c = ref.PoseidonConfig(fx["full_rounds"], fx["partial_rounds"], fx["alpha"],
                                   [[int(v) for v in r] for r in fx["mds"]],
                                   [[int(v) for v in r] for r in fx["ark"]], fx["rate"], fx["capacity"],
                                   ref.BLS12_381_FR)
        elif name.startswith("bls_default_r"):
            rate = int(name[len("bls_default_r"):].split("_")[0])
            c = ref.get_default_poseidon_parameters(rate, name.endswith("_w"))
        elif name == "bn254_t2":
            ark, mds = ref.find_poseidon_ark_and_mds(ref.BN254_FR, 254, 1, 8, 56, 0)
            c = ref.PoseidonConfig(8, 56, 5, mds, ark, 1, 1, ref.BN254_FR)
        elif name == "bn254_t4_cap2":   # capacity 2: exercises capacity-first lane order
            ark, mds = ref.find_poseidon_ark_and_mds(ref.BN254_FR, 254, 3, 8, 56, 0)
            c = ref.PoseidonConfig(8, 56, 5, mds, ark, 2, 2, ref.BN254_FR)
        elif name == "bn254_odd_rf":    # odd full_rounds: extra full round at the end (mod.rs:96)
            ark, mds = ref.find_poseidon_ark_and_mds(ref.BN254_FR, 254, 2, 7, 5, 0)
            c = ref.PoseidonConfig(7, 5, 3, mds, ark, 2, 1, ref.BN254_FR)
        elif name == "bn254_full_only":  # no partial rounds at all
            ark, mds = ref.find_poseidon_ark_and_mds(ref.BN254_FR, 254, 2, 6, 0, 0)
            c = ref.PoseidonConfig(6, 0, 5, mds, ark, 2, 1, ref.BN254_FR)
        elif name == "bn254_rf1":        # full_rounds / 2 == 0: the optimised form does not apply
            ark, mds = ref.find_poseidon_ark_and_mds(ref.BN254_FR, 254, 2, 1, 4, 0)
            c = ref.PoseidonConfig(1, 4, 5, mds, ark, 2, 1, ref.BN254_FR)
        elif name == "bn254_rp1":        # one partial round: it is also the last one (variant 5 runs it as T rows)
            ark, mds = ref.find_poseidon_ark_and_mds(ref.BN254_FR, 254, 2, 4, 1, 0)
            c = ref.PoseidonConfig(4, 1, 5, mds, ark, 2, 1, ref.BN254_FR)
        elif name == "bn254_rp2_alpha17":  # two partial rounds, alpha = 17 on BN254
            ark, mds = ref.find_poseidon_ark_and_mds(ref.BN254_FR, 254, 2, 6, 2, 0)
            c = ref.PoseidonConfig(6, 2, 17, mds, ark, 2, 1, ref.BN254_FR)
        elif name == "bls_t3_alpha5":    # BLS12-381 with alpha = 5 (default table, rate 3 has alpha 5 but t = 4)
            ark, mds = ref.find_poseidon_ark_and_mds(ref.BLS12_381_FR, 255, 2, 8, 56, 0)
            c = ref.PoseidonConfig(8, 56, 5, mds, ark, 2, 1, ref.BLS12_381_FR)
        else:
            raise KeyError(name)
        _CFG_CACHE[name] = c
    return _CFG_CACHE[name]


def field_name(ref, cfg):
    return "bn254" if cfg.p == ref.BN254_FR else "bls12_381"


def field_id(ref, cfg):
    return 0 if cfg.p == ref.BN254_FR else 1
