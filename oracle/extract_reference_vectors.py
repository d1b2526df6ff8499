#!/usr/bin/env python3
"""Extract the known-answer values and the literal parameter fixture that the reference's own
tests hold, into small JSON fixtures under tests/golden/.  TEST INFRASTRUCTURE ONLY.

Reads /root/reference (present only in the build container, never on the GPU box); the JSON
outputs are committed and are what the tests read.  Only *numbers* are extracted (decimal
literals inside `MontFp!("...")` / `F::from_str("...")`), no code.

  reference_kats.json            <- poseidon/grain_lfsr.rs:197-213, poseidon/traits.rs:163-358,
                                    poseidon/mod.rs:376-399          (33 values)
  bls12_381_fixture_params.json  <- poseidon/tests.rs:120-820 (`poseidon_parameters_for_test`)

Usage: python oracle/extract_reference_vectors.py [/root/reference]
"""
import json
import os
import re
import sys

REF = sys.argv[1] if len(sys.argv) > 1 else "/root/reference"
OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests", "golden")

DEC = re.compile(r'"(\d{20,})"')


def read(rel):
    with open(os.path.join(REF, "src", rel)) as f:
        return f.read()


def main():
    os.makedirs(OUT, exist_ok=True)

    # --- grain_lfsr.rs test: first two rejection samples of Grain(false,255,3,8,31)
    lfsr_src = read("poseidon/grain_lfsr.rs")
    lfsr_src = lfsr_src[lfsr_src.index("fn test_grain_lfsr_consistency"):]
    lfsr = DEC.findall(lfsr_src)
    assert len(lfsr) == 2

    # --- traits.rs test: (ark[0][0], mds[0][0]) for rates 2..8 x {constraints, weights}
    tr = read("poseidon/traits.rs")
    tr = tr[tr.index("fn bls12_381_fr_poseidon_default_parameters_test"):]
    # each block: `let <kind>_rate_<r> = ...get_default_poseidon_parameters(<r>, <bool>)`
    blocks = re.split(r"let (constraints|weights)_rate_(\d) =", tr)[1:]
    params = []
    for i in range(0, len(blocks), 3):
        kind, rate, body = blocks[i], int(blocks[i + 1]), blocks[i + 2]
        m = re.search(r"get_default_poseidon_parameters\((\d), (true|false)\)", body)
        assert int(m.group(1)) == rate and (m.group(2) == "true") == (kind == "weights")
        ark00 = DEC.search(body[body.index(".ark[0][0]"):]).group(1)
        mds00 = DEC.search(body[body.index(".mds[0][0]"):]).group(1)
        params.append({"rate": rate, "optimized_for_weights": kind == "weights",
                       "ark00": ark00, "mds00": mds00})
    assert len(params) == 14

    # --- mod.rs test: absorb [0,1,2], squeeze 3 with default params (2,false)
    md = read("poseidon/mod.rs")
    md = md[md.index("fn test_poseidon_sponge_consistency"):]
    sponge = DEC.findall(md)
    assert len(sponge) == 3

    kats = {
        "_source": "arkworks-rs/sponge reference tests; extracted by oracle/extract_reference_vectors.py",
        "field": "bls12_381",
        "grain_lfsr": {"args": [False, 255, 3, 8, 31], "rejection_samples": lfsr,
                       "ref": "poseidon/grain_lfsr.rs:197-213"},
        "default_parameters": {"entries": params, "ref": "poseidon/traits.rs:163-358"},
        "sponge_consistency": {"params": {"rate": 2, "optimized_for_weights": False},
                               "absorb": ["0", "1", "2"], "squeeze_native": sponge,
                               "ref": "poseidon/mod.rs:376-399"},
    }
    with open(os.path.join(OUT, "reference_kats.json"), "w") as f:
        json.dump(kats, f, indent=1)
        f.write("\n")

    # --- tests.rs fixture: alpha=17, 3x3 mds, 37x3 ark, RF=8, RP=29, rate 2, capacity 1
    ts = read("poseidon/tests.rs")
    ts = ts[ts.index("fn poseidon_parameters_for_test"):]
    mds_src = ts[ts.index("let mds"):ts.index("let ark")]
    ark_src = ts[ts.index("let ark"):ts.index("let full_rounds")]
    alpha = int(re.search(r"let alpha = (\d+);", ts).group(1))
    full_rounds = int(re.search(r"let full_rounds = (\d+);", ts).group(1))
    total_rounds = int(re.search(r"let total_rounds = (\d+);", ts).group(1))
    capacity = int(re.search(r"let capacity = (\d+);", ts).group(1))
    rate = int(re.search(r"let rate = (\d+);", ts).group(1))
    mds_flat = DEC.findall(mds_src)
    ark_flat = DEC.findall(ark_src)
    t = rate + capacity
    assert len(mds_flat) == t * t and len(ark_flat) == total_rounds * t
    fixture = {
        "_source": "poseidon/tests.rs:120-820 poseidon_parameters_for_test; extracted by "
                   "oracle/extract_reference_vectors.py",
        "field": "bls12_381", "alpha": alpha, "full_rounds": full_rounds,
        "partial_rounds": total_rounds - full_rounds, "rate": rate, "capacity": capacity,
        "mds": [mds_flat[i * t:(i + 1) * t] for i in range(t)],
        "ark": [ark_flat[i * t:(i + 1) * t] for i in range(total_rounds)],
    }
    with open(os.path.join(OUT, "bls12_381_fixture_params.json"), "w") as f:
        json.dump(fixture, f, indent=1)
        f.write("\n")
    print("wrote", os.path.normpath(OUT))


if __name__ == "__main__":
    main()
