"""ctypes loader for libposeidon_b200.so -- the C ABI declared in include/poseidon_b200.h.

There is no fallback: if the shared library has not been built (`./build.sh` or
`__graft_entry__.build()`), importing this module raises, and if no CUDA device is present every
context creation fails with PSD_ERR_CUDA.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
# POSEIDON_B200_LIB overrides the library path (tuning builds); the default is the in-tree build.
LIB_PATH = os.environ.get("POSEIDON_B200_LIB") or os.path.join(_HERE, "lib", "libposeidon_b200.so")

PSD_OK, PSD_ERR_ARG, PSD_ERR_SHAPE, PSD_ERR_NONCANONICAL, PSD_ERR_CUDA, PSD_ERR_UNSUPPORTED, PSD_ERR_NOMEM = range(7)
FIELD_BN254_FR, FIELD_BLS12_381_FR = 0, 1
REPR_CANONICAL, REPR_MONTGOMERY = 0, 1
LAYOUT_AOS, LAYOUT_SOA = 0, 1
MEM_HOST, MEM_DEVICE = 0, 1
MODE_ABSORBING, MODE_SQUEEZING = 0, 1


class PsdParams(C.Structure):
    """struct psd_params (mirrors PoseidonConfig<F>, reference src/poseidon/mod.rs:23-42)."""
    _fields_ = [
        ("field", C.c_int32),
        ("full_rounds", C.c_uint32),
        ("partial_rounds", C.c_uint32),
        ("alpha", C.c_uint64),
        ("rate", C.c_uint32),
        ("capacity", C.c_uint32),
        ("ark", C.c_void_p),
        ("mds", C.c_void_p),
        ("repr", C.c_int32),
    ]


class PsdCtxInfo(C.Structure):
    """struct psd_ctx_info (psd_ctx_describe)."""
    _fields_ = [
        ("variant", C.c_int32),
        ("constants_home", C.c_int32),
        ("image_bytes", C.c_uint32),
        ("image_shared_by", C.c_uint32),
        ("kernel_name", C.c_char * 96),
    ]


CONSTANTS_BANK, CONSTANTS_SHARED = 0, 1


class PoseidonB200Error(RuntimeError):
    def __init__(self, status: int, message: str):
        super().__init__(f"{status_name(status)}: {message}")
        self.status = status


if not os.path.exists(LIB_PATH):
    raise ImportError(
        f"{LIB_PATH} not found: build the CUDA library first (./build.sh); sponge_b200 has no CPU path")

lib = C.CDLL(LIB_PATH)

_vp, _sz, _i, _u32, _u64 = C.c_void_p, C.c_size_t, C.c_int, C.c_uint32, C.c_uint64
_pp = C.POINTER(C.c_void_p)

_PROTOS = {
    "psd_version": (C.c_char_p, []),
    "psd_last_error_string": (C.c_char_p, []),
    "psd_status_string": (C.c_char_p, [_i]),
    "psd_device_count": (_i, []),
    "psd_ctx_create": (_i, [C.POINTER(PsdParams), _i, _pp]),
    "psd_ctx_destroy": (None, [_vp]),
    "psd_ctx_get_params": (_i, [_vp, C.POINTER(PsdParams)]),
    "psd_find_poseidon_ark_and_mds": (_i, [C.c_int32, _u64, _u32, _u32, _u32, _u32, _vp, _vp]),
    "psd_default_params": (_i, [C.c_int32, _u32, _i, C.POINTER(PsdParams), _vp, _vp]),
    "psd_to_montgomery": (_i, [_vp, _vp, _sz, _i, _vp]),
    "psd_from_montgomery": (_i, [_vp, _vp, _sz, _i, _vp]),
    "psd_permute_batch": (_i, [_vp, _vp, _sz, _i, _i, _i, _vp]),
    "psd_absorb_squeeze_batch": (_i, [_vp, _vp, _sz, _u32, _u32, _vp, _i, _i, _i, _vp]),
    "psd_absorb_squeeze_bytes_batch": (_i, [_vp, _vp, _sz, _u32, _u32, _vp, _i, _i, _i, _vp]),
    "psd_absorb_squeeze_ragged_batch": (_i, [_vp, _vp, _vp, _sz, _u32, _vp, _i, _i, _vp]),
    "psd_merkle_root": (_i, [_vp, _vp, _sz, _vp, _i, _i, _vp]),
    "psd_merkle_tree": (_i, [_vp, _vp, _sz, _vp, _i, _i, _vp]),
    "psd_merkle_path": (_i, [_vp, _vp, _sz, _sz, _vp]),
    "psd_compress_pairs": (_i, [_vp, _vp, _sz, _vp, _i, _i, _vp]),
    "psd_merkle_root_noalloc": (_i, [_vp, _vp, _sz, _vp, _i, _vp]),
    "psd_shard_range": (None, [_sz, _i, _i, C.POINTER(_sz), C.POINTER(_sz)]),
    "psd_permute_batch_multi": (_i, [_pp, _i, _vp, _sz, _i, _i]),
    "psd_absorb_squeeze_batch_multi": (_i, [_pp, _i, _vp, _sz, _u32, _u32, _vp, _i, _i]),
    "psd_merkle_root_multi": (_i, [_pp, _i, _pp, _sz, _vp, _i, _i, _pp]),
    "psd_batch_new": (_i, [_vp, _sz, _pp]),
    "psd_batch_clone": (_i, [_vp, _pp]),
    "psd_batch_free": (None, [_vp]),
    "psd_batch_absorb": (_i, [_vp, _vp, _u32, _i, _i, _i]),
    "psd_batch_squeeze_native": (_i, [_vp, _u32, _vp, _i, _i, _i]),
    "psd_batch_squeeze_bytes": (_i, [_vp, _u32, _vp, _i]),
    "psd_batch_get_state": (_i, [_vp, _vp, _i, _i, _i, C.POINTER(_i), C.POINTER(_u32)]),
    "psd_batch_set_state": (_i, [_vp, _vp, _i, _i, _i, _i, _u32]),
    "psd_batch_size": (_sz, [_vp]),
    "psd_batch_set_stream": (_i, [_vp, _vp]),
    "psd_sponge_new": (_i, [_vp, _pp]),
    "psd_sponge_clone": (_i, [_vp, _pp]),
    "psd_sponge_free": (None, [_vp]),
    "psd_sponge_absorb": (_i, [_vp, _vp, _sz, _i]),
    "psd_sponge_absorb_bytes": (_i, [_vp, _vp, _sz]),
    "psd_sponge_squeeze_native": (_i, [_vp, _sz, _vp, _i]),
    "psd_sponge_squeeze_bytes": (_i, [_vp, _sz, _vp]),
    "psd_sponge_squeeze_bits": (_i, [_vp, _sz, _vp]),
    "psd_sponge_squeeze_field_elements": (_i, [_vp, C.c_int32, _sz, _vp]),
    "psd_sponge_fork": (_i, [_vp, _vp, _sz, _pp]),
    "psd_sponge_get_state": (_i, [_vp, _vp, _i, C.POINTER(_i), C.POINTER(_u32)]),
    "psd_sponge_set_state": (_i, [_vp, _vp, _i, _i, _u32]),
    "psd_host_alloc": (_i, [_sz, _pp]),
    "psd_host_free": (None, [_vp]),
    "psd_device_alloc": (_i, [_vp, _sz, _pp]),
    "psd_device_free": (None, [_vp, _vp]),
    "psd_memcpy_h2d": (_i, [_vp, _vp, _vp, _sz, _vp]),
    "psd_memcpy_d2h": (_i, [_vp, _vp, _vp, _sz, _vp]),
    "psd_stream_synchronize": (_i, [_vp, _vp]),
    "psd_synth_fill": (_i, [_vp, _vp, _sz, _u64, _u64, _i, _vp]),
    "psd_kernel_launch_count": (_u64, []),
    "psd_imad_peak": (_i, [_vp, _i, C.POINTER(C.c_double), C.POINTER(C.c_double)]),
    "psd_ctx_set_variant": (_i, [_vp, _i]),
    "psd_ctx_active_variant": (_i, [_vp, C.POINTER(C.c_int)]),
    "psd_ctx_describe": (_i, [_vp, C.POINTER(PsdCtxInfo)]),
}

for _name, (_res, _args) in _PROTOS.items():
    _fn = getattr(lib, _name)
    _fn.restype = _res
    _fn.argtypes = _args


def status_name(status: int) -> str:
    return lib.psd_status_string(int(status)).decode()


def check(status: int) -> None:
    if status != PSD_OK:
        raise PoseidonB200Error(status, lib.psd_last_error_string().decode())
