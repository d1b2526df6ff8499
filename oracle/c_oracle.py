"""ctypes wrapper of oracle/liboracle.so (the C restatement).  TEST INFRASTRUCTURE ONLY: imported
by tests/, __graft_entry__.smoke() and bench.py's CPU-baseline legs, never by sponge_b200."""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
# PSD_SANITIZE=1 (tests/test_sanitizers.py) selects an ASan + UBSan build of the same source
_SAN = os.environ.get("PSD_SANITIZE") == "1"
LIB_PATH = os.path.join(_HERE, "liboracle_san.so" if _SAN else "liboracle.so")
SRC_PATH = os.path.join(_HERE, "poseidon_oracle.c")


def build(force: bool = False) -> str:
    if force or not os.path.exists(LIB_PATH) or os.path.getmtime(LIB_PATH) < os.path.getmtime(SRC_PATH):
        flags = ["-O1", "-g", "-fsanitize=address,undefined", "-fno-sanitize-recover=all"] if _SAN else ["-O3"]
        subprocess.check_call(["gcc"] + flags + ["-fopenmp", "-shared", "-fPIC", "-o", LIB_PATH, SRC_PATH])
    return LIB_PATH


_lib = None


def lib():
    global _lib
    if _lib is None:
        build()
        l = C.CDLL(LIB_PATH)
        vp, sz, i, u32, u64 = C.c_void_p, C.c_size_t, C.c_int, C.c_uint32, C.c_uint64
        l.orc_create.restype = vp
        l.orc_create.argtypes = [i, u32, u32, u64, u32, u32, vp, vp]
        l.orc_destroy.argtypes = [vp]
        l.orc_to_mont.argtypes = [vp, vp, sz]
        l.orc_from_mont.argtypes = [vp, vp, sz]
        l.orc_max_threads.restype = i
        l.orc_permute_batch.argtypes = [vp, vp, sz, i]
        l.orc_absorb_squeeze_batch.argtypes = [vp, vp, sz, u32, u32, vp, i]
        l.orc_absorb_squeeze_bytes_batch.argtypes = [vp, vp, sz, u32, u32, vp, i]
        l.orc_merkle_root.argtypes = [vp, vp, sz, vp, i]
        l.orc_sponge_new.restype = vp
        l.orc_sponge_new.argtypes = [vp]
        l.orc_sponge_free.argtypes = [vp]
        l.orc_sponge_absorb.argtypes = [vp, vp, sz]
        l.orc_sponge_squeeze.argtypes = [vp, vp, sz]
        l.orc_sponge_state.argtypes = [vp, vp, C.POINTER(i), C.POINTER(u32)]
        l.orc_synth_fill.argtypes = [vp, vp, sz, u64, u64, i, i]
        _lib = l
    return _lib


def _limbs(values):
    vals = list(values)
    return np.frombuffer(b"".join(int(v).to_bytes(32, "little") for v in vals), dtype=np.uint64).reshape(len(vals), 4).copy()


class COracle:
    """One parameter set.  `cfg` is any object with full_rounds, partial_rounds, alpha, mds, ark, rate,
    capacity (canonical ints) and a field id (0 = BN254 Fr, 1 = BLS12-381 Fr)."""

    def __init__(self, cfg, field_id: int):
        self.t = cfg.rate + cfg.capacity
        self.rate, self.capacity = cfg.rate, cfg.capacity
        ark = _limbs(v for row in cfg.ark for v in row)
        mds = _limbs(v for row in cfg.mds for v in row)
        self.h = lib().orc_create(field_id, cfg.full_rounds, cfg.partial_rounds, cfg.alpha, cfg.rate, cfg.capacity,
                                  ark.ctypes.data, mds.ctypes.data)
        if not self.h:
            raise ValueError("orc_create rejected the parameters")

    def __del__(self):
        if getattr(self, "h", None):
            lib().orc_destroy(self.h)
            self.h = None

    def to_mont(self, a: np.ndarray) -> np.ndarray:
        lib().orc_to_mont(self.h, a.ctypes.data, a.size // 4)
        return a

    def from_mont(self, a: np.ndarray) -> np.ndarray:
        lib().orc_from_mont(self.h, a.ctypes.data, a.size // 4)
        return a

    def permute_batch(self, states: np.ndarray, threads: int = 0) -> np.ndarray:
        lib().orc_permute_batch(self.h, states.ctypes.data, states.size // (4 * self.t), threads)
        return states

    def absorb_squeeze_batch(self, inputs: np.ndarray, n: int, n_absorb: int, n_squeeze: int, threads: int = 0):
        out = np.zeros((n, n_squeeze, 4), dtype=np.uint64)
        lib().orc_absorb_squeeze_batch(self.h, inputs.ctypes.data, n, n_absorb, n_squeeze, out.ctypes.data, threads)
        return out

    def absorb_squeeze_bytes_batch(self, inputs: np.ndarray, n: int, n_absorb: int, n_bytes: int, threads: int = 0):
        out = np.zeros((n, n_bytes), dtype=np.uint8)
        lib().orc_absorb_squeeze_bytes_batch(self.h, inputs.ctypes.data, n, n_absorb, n_bytes, out.ctypes.data, threads)
        return out

    def merkle_root(self, leaves: np.ndarray, threads: int = 0) -> np.ndarray:
        root = np.zeros(4, dtype=np.uint64)
        lib().orc_merkle_root(self.h, leaves.ctypes.data, leaves.size // 4, root.ctypes.data, threads)
        return root

    def synth(self, n: int, seed: int, first: int = 0, montgomery: bool = True, threads: int = 0) -> np.ndarray:
        out = np.zeros((n, 4), dtype=np.uint64)
        lib().orc_synth_fill(self.h, out.ctypes.data, n, seed, first, int(montgomery), threads)
        return out

    def max_threads(self) -> int:
        return lib().orc_max_threads()


class COracleSponge:
    def __init__(self, orc: COracle):
        self.orc = orc
        self.h = lib().orc_sponge_new(orc.h)

    def __del__(self):
        if getattr(self, "h", None):
            lib().orc_sponge_free(self.h)
            self.h = None

    def absorb(self, elems_mont: np.ndarray) -> None:
        lib().orc_sponge_absorb(self.h, elems_mont.ctypes.data, elems_mont.size // 4)

    def squeeze(self, n: int) -> np.ndarray:
        out = np.zeros((max(n, 1), 4), dtype=np.uint64)
        lib().orc_sponge_squeeze(self.h, out.ctypes.data, n)
        return out[:n]

    def state(self):
        st = np.zeros((self.orc.t, 4), dtype=np.uint64)
        mode, idx = C.c_int(), C.c_uint32()
        lib().orc_sponge_state(self.h, st.ctypes.data, C.byref(mode), C.byref(idx))
        return st, mode.value, idx.value
