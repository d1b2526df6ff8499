"""AddressSanitizer + UndefinedBehaviorSanitizer pass over the two pieces of C/C++ test infrastructure
that every parity claim leans on: the C oracle (oracle/poseidon_oracle.c) and the host model of the device
arithmetic (tests/emu/fp_emu.cpp over sponge_b200/csrc/*.cuh).  The oracle and host-model tests are
re-run in a child interpreter with PSD_SANITIZE=1, which makes their loaders build and load
`-fsanitize=address,undefined -fno-sanitize-recover=all` variants; libasan is preloaded because the
interpreter itself is not instrumented.  Any out-of-bounds access, signed overflow or misaligned access
aborts the child.  CPU only (compute-sanitizer is not available on the GPU pool)."""
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _libasan():
    try:
        p = subprocess.check_output(["gcc", "-print-file-name=libasan.so"], text=True).strip()
    except (OSError, subprocess.CalledProcessError):
        return None
    return p if os.path.isabs(p) and os.path.exists(p) else None


@pytest.mark.skipif(os.environ.get("PSD_SANITIZE") == "1", reason="already inside the sanitizer run")
def test_oracle_and_host_model_under_asan_ubsan():
    asan = _libasan()
    if asan is None:
        pytest.skip("gcc has no libasan on this machine")
    env = dict(os.environ, PSD_SANITIZE="1", LD_PRELOAD=asan,
               ASAN_OPTIONS="detect_leaks=0:abort_on_error=1", UBSAN_OPTIONS="halt_on_error=1:print_stacktrace=1")
    sel = ["tests/test_oracle.py::test_c_oracle_matches_python_oracle", "tests/test_oracle.py::test_c_oracle_bytes_merkle_synth",
           "tests/test_host_logic.py::test_device_arithmetic_host_model",
           "tests/test_host_logic.py::test_permutation_variants_host_model",
           "tests/test_host_logic.py::test_lazy_sbox_host_model", "tests/test_host_logic.py::test_mont_dot_host_model",
           "tests/test_host_logic.py::test_squaring_and_wide_window_host_model",
           "tests/test_host_logic.py::test_fp52_primitives_host_model"]
    r = subprocess.run([sys.executable, "-m", "pytest", "-x", "-q", "-p", "no:cacheprovider"] + sel, cwd=ROOT, env=env,
                       capture_output=True, text=True, timeout=1500)
    assert r.returncode == 0, r.stdout[-4000:] + r.stderr[-4000:]
