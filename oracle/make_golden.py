#!/usr/bin/env python3
"""Generate tests/golden/derived_vectors.json from the KAT-gated Python oracle.  TEST INFRASTRUCTURE.

These values are NOT pinned by the reference (it has no BN254 type, and its literal test fixture
ships without expected outputs); they are produced by oracle/poseidon_ref.py -- the code that
reproduces all 33 reference KATs -- and serve as regression anchors for the C oracle and the GPU.
The one externally known value is circomlib's poseidon([1,2]) (bn254_t3_permute_0_1_2[0]).

Usage: python oracle/make_golden.py
"""
import json
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
from oracle import poseidon_ref as R  # noqa: E402

GOLDEN = os.path.join(HERE, "..", "tests", "golden")


def fixture_cfg():
    fx = json.load(open(os.path.join(GOLDEN, "bls12_381_fixture_params.json")))
    return R.PoseidonConfig(fx["full_rounds"], fx["partial_rounds"], fx["alpha"],
                            [[int(v) for v in r] for r in fx["mds"]], [[int(v) for v in r] for r in fx["ark"]],
                            fx["rate"], fx["capacity"], R.BLS12_381_FR)


def main():
    out = {"_source": "oracle/make_golden.py (oracle/poseidon_ref.py, KAT-gated); values unpinned by the reference"}
    bn3 = R.bn254_config(2, 8, 57)
    bn5 = R.bn254_config(4, 8, 60)
    fx = fixture_cfg()
    dflt = R.get_default_poseidon_parameters(2, False)
    h = lambda v: "0x%064x" % v

    out["bn254_t3_ark00"] = h(bn3.ark[0][0])
    out["bn254_t3_mds00"] = h(bn3.mds[0][0])
    out["bn254_t3_permute_0_1_2"] = [h(v) for v in R.permute(bn3, [0, 1, 2])]
    out["bn254_t3_absorb_1_2_squeeze_1"] = h(R.hash_absorb_squeeze(bn3, [1, 2], 1)[0])
    out["bn254_t3_absorb_1_2_squeeze_1_montgomery_limbs64"] = [
        "0x%016x" % l for l in R.int_to_limbs64(R.to_mont(R.hash_absorb_squeeze(bn3, [1, 2], 1)[0], R.BN254_FR))]
    out["bn254_t5_ark00"] = h(bn5.ark[0][0])
    out["bn254_t5_mds00"] = h(bn5.mds[0][0])
    out["bn254_t5_absorb_1_2_3_4_squeeze_1"] = h(R.hash_absorb_squeeze(bn5, [1, 2, 3, 4], 1)[0])
    out["bls_fixture_absorb_1_2_squeeze_1"] = h(R.hash_absorb_squeeze(fx, [1, 2], 1)[0])
    s = R.PoseidonSponge(fx)
    s.absorb(list(range(1024)))
    out["bls_fixture_absorb_0_1023_squeeze_bytes_32"] = s.squeeze_bytes(32).hex()
    out["bls_fixture_absorb_0_1023_permutes"] = s.n_permutes
    out["bn254_t3_merkle_leaves_0_7"] = h(R.merkle_root(bn3, list(range(8))))
    out["bn254_t3_merkle_leaves_1_1024"] = h(R.merkle_root(bn3, list(range(1, 1025))))
    # quirk Q1 (mod.rs:175) as numbers: absorb([7]); squeeze(1); squeeze(2) at rate 2
    s = R.PoseidonSponge(dflt)
    s.absorb([7])
    first = s.squeeze_native_field_elements(1)
    second = s.squeeze_native_field_elements(2)
    out["bls_default_q1"] = {"first": [str(v) for v in first], "second": [str(v) for v in second],
                             "permutes": s.n_permutes}
    # bytes absorb + fork (absorb.rs:122-140, lib.rs:149-157)
    s = R.PoseidonSponge(fx)
    s.absorb_bytes(bytes(range(100)))
    out["bls_fixture_absorb_bytes_0_99_squeeze_2"] = [str(v) for v in s.squeeze_native_field_elements(2)]
    s = R.PoseidonSponge(fx)
    s.absorb([5])
    f = s.fork(b"domain-A")
    out["bls_fixture_absorb_5_fork_domainA_squeeze_2"] = [str(v) for v in f.squeeze_native_field_elements(2)]
    s = R.PoseidonSponge(fx)
    s.absorb([1, 2, 3])
    out["bls_fixture_absorb_1_2_3_squeeze_bits_300"] = "".join("1" if b else "0" for b in s.squeeze_bits(300))
    s = R.PoseidonSponge(fx)
    s.absorb([1, 2, 3])
    out["bls_fixture_absorb_1_2_3_squeeze_bn254_3"] = [str(v) for v in s.squeeze_field_elements(3, R.BN254_FR)]
    with open(os.path.join(GOLDEN, "derived_vectors.json"), "w") as fo:
        json.dump(out, fo, indent=1)
        fo.write("\n")
    print("wrote derived_vectors.json")


if __name__ == "__main__":
    main()
