"""Host-side mirror of ark-sponge's Poseidon interface over the B200 C ABI.

Same names, argument meaning and error behaviour as the reference (paths relative to
/root/reference/src):

  PoseidonConfig / PoseidonParameters   poseidon/mod.rs:23-42, ::new 187-213 (mds BEFORE ark)
  PoseidonSponge                        poseidon/mod.rs:44-60, CryptographicSponge impl 216-318,
                                        FieldBasedCryptographicSponge 320-342, SpongeExt 344-367
  DuplexSpongeMode, FieldElementSize    lib.rs:198-210, 35-58
  find_poseidon_ark_and_mds,
  get_default_poseidon_parameters       poseidon/traits.rs:59-155

plus the batch entry points the build adds (PoseidonContext.permute_batch /
absorb_squeeze_batch / merkle_root, SpongeBatch).  Field elements are Python ints (canonical
values) on the single-sponge surface and numpy uint64 arrays of shape (..., 4) -- the memory
image of `ark_ff::Fp<MontBackend<_,4>,4>` -- on the batch surface.

Every computation happens in libposeidon_b200.so on the GPU; nothing here computes a permutation.
"""
from __future__ import annotations

import ctypes as C
import weakref
from dataclasses import dataclass
from typing import Iterable, List, Optional, Sequence, Tuple, Union

import numpy as np

from . import _lib as L
from ._lib import PoseidonB200Error, check, lib

FIELDS = {"bn254": L.FIELD_BN254_FR, "bls12_381": L.FIELD_BLS12_381_FR}
MODULUS = {
    "bn254": 21888242871839275222246405745257275088548364400416034343698204186575808495617,
    "bls12_381": 52435875175126190479447740508185965837690552500527637822603658699938581184513,
}
MODULUS_BIT_SIZE = {"bn254": 254, "bls12_381": 255}


# ---------------------------------------------------------------------------------------------
# element <-> limb helpers
# ---------------------------------------------------------------------------------------------
def ints_to_limbs(values: Iterable[int]) -> np.ndarray:
    """Python ints -> (n, 4) uint64 little-endian limbs."""
    vals = list(values)
    buf = b"".join(int(v).to_bytes(32, "little") for v in vals)
    return np.frombuffer(buf, dtype=np.uint64).reshape(len(vals), 4).copy()


def limbs_to_ints(arr: np.ndarray) -> List[int]:
    a = np.ascontiguousarray(arr, dtype=np.uint64).reshape(-1, 4)
    raw = a.tobytes()
    return [int.from_bytes(raw[i:i + 32], "little") for i in range(0, len(raw), 32)]


def _ptr(a: Optional[np.ndarray]) -> Optional[int]:
    return None if a is None else a.ctypes.data


# ---------------------------------------------------------------------------------------------
# lib.rs enums
# ---------------------------------------------------------------------------------------------
@dataclass(frozen=True)
class DuplexSpongeMode:
    """lib.rs:198-210.  kind is 'Absorbing' (index = next_absorb_index) or 'Squeezing'."""
    kind: str
    index: int

    @staticmethod
    def Absorbing(next_absorb_index: int) -> "DuplexSpongeMode":
        return DuplexSpongeMode("Absorbing", next_absorb_index)

    @staticmethod
    def Squeezing(next_squeeze_index: int) -> "DuplexSpongeMode":
        return DuplexSpongeMode("Squeezing", next_squeeze_index)


@dataclass(frozen=True)
class FieldElementSize:
    """lib.rs:35-58.  truncated is None for Full, else the bit count of Truncated(n)."""
    truncated: Optional[int] = None

    @staticmethod
    def Full() -> "FieldElementSize":
        return FieldElementSize(None)

    @staticmethod
    def Truncated(n: int) -> "FieldElementSize":
        return FieldElementSize(n)

    def num_bits(self, field: str) -> int:
        # lib.rs:45-52: Truncated(n) is validated, then MODULUS_BIT_SIZE-1 is returned regardless
        if self.truncated is not None and self.truncated > MODULUS_BIT_SIZE[field]:
            raise ValueError("num_bits is greater than the capacity of the field.")
        return MODULUS_BIT_SIZE[field] - 1


# ---------------------------------------------------------------------------------------------
# parameters
# ---------------------------------------------------------------------------------------------
class PoseidonConfig:
    """poseidon/mod.rs:23-42.  ark[round][lane], mds[row][col]; canonical Python ints."""

    def __init__(self, full_rounds: int, partial_rounds: int, alpha: int, mds: Sequence[Sequence[int]],
                 ark: Sequence[Sequence[int]], rate: int, capacity: int, field: str = "bls12_381"):
        if field not in FIELDS:
            raise ValueError(f"unknown field {field!r}")
        # shape asserts of PoseidonConfig::new, mod.rs:196-203
        t = rate + capacity
        assert len(ark) == full_rounds + partial_rounds, "ark.len() != full_rounds + partial_rounds"
        for row in ark:
            assert len(row) == t, "ark row length != rate + capacity"
        assert len(mds) == t, "mds.len() != rate + capacity"
        for row in mds:
            assert len(row) == t, "mds row length != rate + capacity"
        self.full_rounds, self.partial_rounds, self.alpha = full_rounds, partial_rounds, alpha
        self.mds = [list(r) for r in mds]
        self.ark = [list(r) for r in ark]
        self.rate, self.capacity, self.field = rate, capacity, field
        self._ctx: Optional["PoseidonContext"] = None

    @classmethod
    def new(cls, full_rounds, partial_rounds, alpha, mds, ark, rate, capacity, field="bls12_381"):
        """Same argument order as the reference: mds before ark (mod.rs:187-195)."""
        return cls(full_rounds, partial_rounds, alpha, mds, ark, rate, capacity, field)

    @property
    def modulus(self) -> int:
        return MODULUS[self.field]

    def context(self, device: int = 0) -> "PoseidonContext":
        """The GPU-resident form of this parameter set (created once, cached)."""
        if self._ctx is None or self._ctx.closed or self._ctx.device != device:
            self._ctx = PoseidonContext(self, device)
        return self._ctx


PoseidonParameters = PoseidonConfig   # name used by ark-sponge 0.3.0 and by BASELINE.json


def find_poseidon_ark_and_mds(field: str, prime_bits: int, rate: int, full_rounds: int, partial_rounds: int,
                              skip_matrices: int) -> Tuple[List[List[int]], List[List[int]]]:
    """poseidon/traits.rs:105-146 (host-only set-up, runs in the C++ library)."""
    t = rate + 1
    ark = np.zeros(((full_rounds + partial_rounds) * t, 4), dtype=np.uint64)
    mds = np.zeros((t * t, 4), dtype=np.uint64)
    check(lib.psd_find_poseidon_ark_and_mds(FIELDS[field], prime_bits, rate, full_rounds, partial_rounds,
                                            skip_matrices, _ptr(ark), _ptr(mds)))
    a, m = limbs_to_ints(ark), limbs_to_ints(mds)
    return ([a[i * t:(i + 1) * t] for i in range(full_rounds + partial_rounds)],
            [m[i * t:(i + 1) * t] for i in range(t)])


def get_default_poseidon_parameters(field: str, rate: int, optimized_for_weights: bool) -> Optional[PoseidonConfig]:
    """PoseidonDefaultConfigField::get_default_poseidon_parameters (traits.rs:59-102); None where the
    reference returns None."""
    p = L.PsdParams()
    ark = np.zeros((73 * 9, 4), dtype=np.uint64)
    mds = np.zeros((81, 4), dtype=np.uint64)
    st = lib.psd_default_params(FIELDS[field], rate, int(optimized_for_weights), C.byref(p), _ptr(ark), _ptr(mds))
    if st == L.PSD_ERR_UNSUPPORTED:
        return None
    check(st)
    t = p.rate + p.capacity
    rounds = p.full_rounds + p.partial_rounds
    a, m = limbs_to_ints(ark[:rounds * t]), limbs_to_ints(mds[:t * t])
    return PoseidonConfig(p.full_rounds, p.partial_rounds, p.alpha, [m[i * t:(i + 1) * t] for i in range(t)],
                          [a[i * t:(i + 1) * t] for i in range(rounds)], p.rate, p.capacity, field)


# ---------------------------------------------------------------------------------------------
# GPU context + batch entry points
# ---------------------------------------------------------------------------------------------
class PoseidonContext:
    """psd_ctx: one parameter set resident on one GPU.  Replaces the per-sponge `parameters.clone()`
    of PoseidonSponge::new (mod.rs:226)."""

    def __init__(self, cfg: PoseidonConfig, device: int = 0):
        self.cfg, self.device = cfg, device
        t = cfg.rate + cfg.capacity
        p = cfg.modulus
        for row in list(cfg.ark) + list(cfg.mds):
            for v in row:
                if not 0 <= v < p:
                    raise PoseidonB200Error(L.PSD_ERR_NONCANONICAL, "parameter value outside [0,p)")
        self._ark = ints_to_limbs(v for row in cfg.ark for v in row)
        self._mds = ints_to_limbs(v for row in cfg.mds for v in row)
        params = L.PsdParams(FIELDS[cfg.field], cfg.full_rounds, cfg.partial_rounds, cfg.alpha, cfg.rate,
                             cfg.capacity, _ptr(self._ark), _ptr(self._mds), L.REPR_CANONICAL)
        h = C.c_void_p()
        check(lib.psd_ctx_create(C.byref(params), device, C.byref(h)))
        self.handle = h
        self.t = t
        self.closed = False
        self._fin = weakref.finalize(self, lib.psd_ctx_destroy, h)

    def close(self) -> None:
        if not self.closed:
            self._fin()
            self.closed = True

    def set_variant(self, variant: int) -> None:
        check(lib.psd_ctx_set_variant(self.handle, variant))

    def active_variant(self) -> int:
        """The kernel variant a launch on this context really runs (requested variant after fallbacks)."""
        v = C.c_int(0)
        check(lib.psd_ctx_active_variant(self.handle, C.byref(v)))
        return v.value

    def describe(self) -> dict:
        """Which kernel a launch on this context runs and where its constants live (psd_ctx_describe)."""
        info = L.PsdCtxInfo()
        check(lib.psd_ctx_describe(self.handle, C.byref(info)))
        return dict(variant=info.variant, constants_home="bank" if info.constants_home == L.CONSTANTS_BANK else "shared",
                    image_bytes=info.image_bytes, image_shared_by=info.image_shared_by,
                    kernel_name=info.kernel_name.decode())

    # -- host-array batch calls ---------------------------------------------------------------
    def permute_batch(self, states: np.ndarray, layout: int = L.LAYOUT_AOS, repr: int = L.REPR_MONTGOMERY) -> np.ndarray:
        """n independent `permute`s (mod.rs:95-118), in place on a uint64 array holding n*t elements."""
        assert states.dtype == np.uint64 and states.flags.c_contiguous and states.size % (4 * self.t) == 0
        n = states.size // (4 * self.t)
        check(lib.psd_permute_batch(self.handle, _ptr(states), n, layout, repr, L.MEM_HOST, None))
        return states

    def absorb_squeeze_batch(self, inputs: np.ndarray, n: int, n_absorb: int, n_squeeze: int,
                             layout: int = L.LAYOUT_AOS, repr: int = L.REPR_MONTGOMERY) -> np.ndarray:
        assert inputs.dtype == np.uint64 and inputs.flags.c_contiguous and inputs.size == n * n_absorb * 4
        shape = (n, n_squeeze, 4) if layout == L.LAYOUT_AOS else (n_squeeze, 2, n, 2)
        out = np.zeros(shape, dtype=np.uint64)
        check(lib.psd_absorb_squeeze_batch(self.handle, _ptr(inputs), n, n_absorb, n_squeeze, _ptr(out), layout, repr,
                                           L.MEM_HOST, None))
        return out

    def absorb_squeeze_ragged_batch(self, inputs: np.ndarray, offsets: np.ndarray, n_squeeze: int,
                                    repr: int = L.REPR_MONTGOMERY) -> np.ndarray:
        """Sponge i absorbs inputs[offsets[i]:offsets[i+1]] (elements) and squeezes n_squeeze elements."""
        offsets = np.ascontiguousarray(offsets, dtype=np.uint64)
        n = offsets.size - 1
        assert inputs.dtype == np.uint64 and inputs.flags.c_contiguous and inputs.size == int(offsets[-1]) * 4
        out = np.zeros((n, n_squeeze, 4), dtype=np.uint64)
        check(lib.psd_absorb_squeeze_ragged_batch(self.handle, _ptr(inputs) if inputs.size else None, _ptr(offsets), n,
                                                  n_squeeze, _ptr(out), repr, L.MEM_HOST, None))
        return out

    def absorb_squeeze_bytes_batch(self, inputs: np.ndarray, n: int, n_absorb: int, n_bytes: int,
                                   layout: int = L.LAYOUT_AOS, repr: int = L.REPR_MONTGOMERY) -> np.ndarray:
        assert inputs.dtype == np.uint64 and inputs.flags.c_contiguous and inputs.size == n * n_absorb * 4
        out = np.zeros((n, n_bytes), dtype=np.uint8)
        check(lib.psd_absorb_squeeze_bytes_batch(self.handle, _ptr(inputs), n, n_absorb, n_bytes, _ptr(out), layout,
                                                 repr, L.MEM_HOST, None))
        return out

    def compress_pairs(self, elems: np.ndarray, repr: int = L.REPR_MONTGOMERY) -> np.ndarray:
        assert elems.dtype == np.uint64 and elems.flags.c_contiguous and elems.size % 8 == 0
        n_pairs = elems.size // 8
        out = np.zeros((n_pairs, 4), dtype=np.uint64)
        check(lib.psd_compress_pairs(self.handle, _ptr(elems), n_pairs, _ptr(out), repr, L.MEM_HOST, None))
        return out

    def merkle_root(self, leaves: np.ndarray, repr: int = L.REPR_MONTGOMERY) -> np.ndarray:
        assert leaves.dtype == np.uint64 and leaves.flags.c_contiguous and leaves.size % 4 == 0
        root = np.zeros(4, dtype=np.uint64)
        check(lib.psd_merkle_root(self.handle, _ptr(leaves), leaves.size // 4, _ptr(root), repr, L.MEM_HOST, None))
        return root

    def merkle_tree(self, leaves: np.ndarray, repr: int = L.REPR_MONTGOMERY) -> np.ndarray:
        """All internal nodes, bottom-up (parents of the leaves first, root last): (n_leaves-1, 4)."""
        assert leaves.dtype == np.uint64 and leaves.flags.c_contiguous and leaves.size % 4 == 0
        n = leaves.size // 4
        nodes = np.zeros((n - 1, 4), dtype=np.uint64)
        check(lib.psd_merkle_tree(self.handle, _ptr(leaves), n, _ptr(nodes), repr, L.MEM_HOST, None))
        return nodes

    @staticmethod
    def merkle_path(leaves: np.ndarray, nodes: np.ndarray, index: int) -> np.ndarray:
        """Authentication path of leaf `index`: the sibling at every level, leaf level first."""
        n = leaves.size // 4
        path = np.zeros((max(n.bit_length() - 1, 1), 4), dtype=np.uint64)
        check(lib.psd_merkle_path(_ptr(leaves), _ptr(nodes), n, index, _ptr(path)))
        return path[: n.bit_length() - 1]

    def to_montgomery(self, elems: np.ndarray) -> np.ndarray:
        check(lib.psd_to_montgomery(self.handle, _ptr(elems), elems.size // 4, L.MEM_HOST, None))
        return elems

    def from_montgomery(self, elems: np.ndarray) -> np.ndarray:
        check(lib.psd_from_montgomery(self.handle, _ptr(elems), elems.size // 4, L.MEM_HOST, None))
        return elems

    # -- device-pointer calls (ints from e.g. torch.Tensor.data_ptr()), asynchronous on `stream` ---
    def permute_batch_dev(self, dptr: int, n: int, layout=L.LAYOUT_AOS, repr=L.REPR_MONTGOMERY, stream: int = 0):
        check(lib.psd_permute_batch(self.handle, dptr, n, layout, repr, L.MEM_DEVICE, stream or None))

    def absorb_squeeze_batch_dev(self, in_ptr: int, n: int, n_absorb: int, n_squeeze: int, out_ptr: int,
                                 layout=L.LAYOUT_AOS, repr=L.REPR_MONTGOMERY, stream: int = 0):
        check(lib.psd_absorb_squeeze_batch(self.handle, in_ptr, n, n_absorb, n_squeeze, out_ptr, layout, repr,
                                           L.MEM_DEVICE, stream or None))

    def absorb_squeeze_bytes_batch_dev(self, in_ptr: int, n: int, n_absorb: int, n_bytes: int, out_ptr: int,
                                       layout=L.LAYOUT_AOS, repr=L.REPR_MONTGOMERY, stream: int = 0):
        check(lib.psd_absorb_squeeze_bytes_batch(self.handle, in_ptr, n, n_absorb, n_bytes, out_ptr, layout, repr,
                                                 L.MEM_DEVICE, stream or None))

    def merkle_root_dev(self, leaves_ptr: int, n_leaves: int, root_ptr: int, repr=L.REPR_MONTGOMERY, stream: int = 0):
        check(lib.psd_merkle_root(self.handle, leaves_ptr, n_leaves, root_ptr, repr, L.MEM_DEVICE, stream or None))

    def synth_fill_dev(self, dptr: int, n: int, seed: int, first_index: int = 0, repr=L.REPR_MONTGOMERY, stream: int = 0):
        check(lib.psd_synth_fill(self.handle, dptr, n, seed, first_index, repr, stream or None))

    def synchronize(self, stream: int = 0) -> None:
        check(lib.psd_stream_synchronize(self.handle, stream or None))

    def imad_peak(self, kind: int) -> Tuple[float, float]:
        rate, ms = C.c_double(), C.c_double()
        check(lib.psd_imad_peak(self.handle, kind, C.byref(rate), C.byref(ms)))
        return rate.value, ms.value


def kernel_launch_count() -> int:
    return int(lib.psd_kernel_launch_count())


def shard_range(n: int, rank: int, world: int) -> Tuple[int, int]:
    """Contiguous shard [lo, hi) of `rank` (psd_shard_range): the first n % world ranks get one extra unit."""
    if not 0 <= rank < world:
        raise ValueError("rank outside [0, world)")
    lo, hi = C.c_size_t(), C.c_size_t()
    lib.psd_shard_range(n, rank, world, C.byref(lo), C.byref(hi))
    return lo.value, hi.value


class MultiContext:
    """One parameter set resident on several GPUs of this process (psd_*_multi entry points): host batches
    are sharded contiguously over the devices with no exchange between them; the Merkle root gathers only
    the per-device subtree roots."""

    def __init__(self, cfg: PoseidonConfig, devices: Sequence[int]):
        if len(devices) == 0:
            raise ValueError("no devices")
        self.cfg, self.devices = cfg, list(devices)
        self.ctxs = [PoseidonContext(cfg, d) for d in self.devices]
        self.t = self.ctxs[0].t
        self._arr = (C.c_void_p * len(self.ctxs))(*[c.handle for c in self.ctxs])

    def close(self) -> None:
        for c in self.ctxs:
            c.close()

    def __len__(self) -> int:
        return len(self.ctxs)

    def permute_batch(self, states: np.ndarray, layout: int = L.LAYOUT_AOS, repr: int = L.REPR_MONTGOMERY) -> np.ndarray:
        assert states.dtype == np.uint64 and states.flags.c_contiguous and states.size % (4 * self.t) == 0
        n = states.size // (4 * self.t)
        check(lib.psd_permute_batch_multi(self._arr, len(self.ctxs), _ptr(states), n, layout, repr))
        return states

    def absorb_squeeze_batch(self, inputs: np.ndarray, n: int, n_absorb: int, n_squeeze: int,
                             layout: int = L.LAYOUT_AOS, repr: int = L.REPR_MONTGOMERY) -> np.ndarray:
        assert inputs.dtype == np.uint64 and inputs.flags.c_contiguous and inputs.size == n * n_absorb * 4
        shape = (n, n_squeeze, 4) if layout == L.LAYOUT_AOS else (n_squeeze, 2, n, 2)
        out = np.zeros(shape, dtype=np.uint64)
        check(lib.psd_absorb_squeeze_batch_multi(self._arr, len(self.ctxs), _ptr(inputs), n, n_absorb, n_squeeze,
                                                 _ptr(out), layout, repr))
        return out

    def merkle_root(self, leaves: np.ndarray, repr: int = L.REPR_MONTGOMERY) -> np.ndarray:
        """Root over one contiguous host array of leaves; device g hashes the g-th n/len(devices) slice."""
        assert leaves.dtype == np.uint64 and leaves.flags.c_contiguous and leaves.size % (4 * len(self.ctxs)) == 0
        per = leaves.size // 4 // len(self.ctxs)
        ptrs = (C.c_void_p * len(self.ctxs))(*[leaves.ctypes.data + g * per * 32 for g in range(len(self.ctxs))])
        root = np.zeros(4, dtype=np.uint64)
        check(lib.psd_merkle_root_multi(self._arr, len(self.ctxs), ptrs, per, _ptr(root), repr, L.MEM_HOST, None))
        return root

    def merkle_root_dev(self, leaf_ptrs: Sequence[int], leaves_per_gpu: int, root_ptr: int,
                        repr: int = L.REPR_MONTGOMERY, streams: Optional[Sequence[int]] = None) -> None:
        """Device-resident form: leaf_ptrs[g] on device g, root_ptr on the first device."""
        ptrs = (C.c_void_p * len(self.ctxs))(*leaf_ptrs)
        st = None if streams is None else (C.c_void_p * len(self.ctxs))(*[s or None for s in streams])
        check(lib.psd_merkle_root_multi(self._arr, len(self.ctxs), ptrs, leaves_per_gpu, root_ptr, repr, L.MEM_DEVICE, st))


class SpongeBatch:
    """n PoseidonSponge values advancing in lock-step (psd_batch): the absorb/squeeze rate
    bookkeeping of mod.rs:121-182, 232-254, 320-342 lifted to a batch; states stay in HBM."""

    def __init__(self, ctx: PoseidonContext, n: int, _handle=None):
        self.ctx, self.n = ctx, n
        if _handle is None:
            _handle = C.c_void_p()
            check(lib.psd_batch_new(ctx.handle, n, C.byref(_handle)))
        self.handle = _handle
        self._fin = weakref.finalize(self, lib.psd_batch_free, _handle)

    def clone(self) -> "SpongeBatch":
        h = C.c_void_p()
        check(lib.psd_batch_clone(self.handle, C.byref(h)))
        return SpongeBatch(self.ctx, self.n, h)

    def free(self) -> None:
        self._fin()

    def set_stream(self, stream: int) -> None:
        check(lib.psd_batch_set_stream(self.handle, stream or None))

    def absorb(self, elems: np.ndarray, k: int, layout=L.LAYOUT_AOS, repr=L.REPR_MONTGOMERY) -> None:
        assert elems.dtype == np.uint64 and elems.flags.c_contiguous and elems.size == self.n * k * 4
        check(lib.psd_batch_absorb(self.handle, _ptr(elems), k, layout, repr, L.MEM_HOST))

    def absorb_dev(self, dptr: int, k: int, layout=L.LAYOUT_AOS, repr=L.REPR_MONTGOMERY) -> None:
        check(lib.psd_batch_absorb(self.handle, dptr, k, layout, repr, L.MEM_DEVICE))

    def squeeze_native(self, k: int, layout=L.LAYOUT_AOS, repr=L.REPR_MONTGOMERY) -> np.ndarray:
        shape = (self.n, k, 4) if layout == L.LAYOUT_AOS else (k, 2, self.n, 2)
        out = np.zeros(shape, dtype=np.uint64)
        check(lib.psd_batch_squeeze_native(self.handle, k, _ptr(out), layout, repr, L.MEM_HOST))
        return out

    def squeeze_bytes(self, n_bytes: int) -> np.ndarray:
        out = np.zeros((self.n, n_bytes), dtype=np.uint8)
        check(lib.psd_batch_squeeze_bytes(self.handle, n_bytes, _ptr(out), L.MEM_HOST))
        return out

    def squeeze_bytes_dev(self, n_bytes: int, out_ptr: int) -> None:
        check(lib.psd_batch_squeeze_bytes(self.handle, n_bytes, out_ptr, L.MEM_DEVICE))

    def get_state(self, layout=L.LAYOUT_AOS, repr=L.REPR_MONTGOMERY) -> Tuple[np.ndarray, DuplexSpongeMode]:
        t = self.ctx.t
        shape = (self.n, t, 4) if layout == L.LAYOUT_AOS else (t, 2, self.n, 2)
        st = np.zeros(shape, dtype=np.uint64)
        mode, idx = C.c_int(), C.c_uint32()
        check(lib.psd_batch_get_state(self.handle, _ptr(st), layout, repr, L.MEM_HOST, C.byref(mode), C.byref(idx)))
        return st, DuplexSpongeMode("Absorbing" if mode.value == 0 else "Squeezing", idx.value)

    def set_state(self, states: np.ndarray, mode: DuplexSpongeMode, layout=L.LAYOUT_AOS, repr=L.REPR_MONTGOMERY) -> None:
        assert states.dtype == np.uint64 and states.flags.c_contiguous and states.size == self.n * self.ctx.t * 4
        check(lib.psd_batch_set_state(self.handle, _ptr(states), layout, repr, L.MEM_HOST,
                                      0 if mode.kind == "Absorbing" else 1, mode.index))


# ---------------------------------------------------------------------------------------------
# The single-sponge drop-in
# ---------------------------------------------------------------------------------------------
AbsorbInput = Union[int, bytes, bytearray, Sequence[int]]


class PoseidonSpongeState:
    """poseidon/mod.rs:344-349: state + mode, no parameters."""

    def __init__(self, state: List[int], mode: DuplexSpongeMode):
        self.state, self.mode = state, mode

    def clone(self) -> "PoseidonSpongeState":
        return PoseidonSpongeState(list(self.state), self.mode)


class PoseidonSponge:
    """CryptographicSponge + FieldBasedCryptographicSponge + SpongeExt for Poseidon
    (lib.rs:105-195, poseidon/mod.rs:216-367), executed on the GPU as a batch of one."""

    def __init__(self, parameters: PoseidonConfig, device: int = 0, _handle=None):
        self.parameters = parameters
        self._ctx = parameters.context(device)
        if _handle is None:
            _handle = C.c_void_p()
            check(lib.psd_sponge_new(self._ctx.handle, C.byref(_handle)))
        self._h = _handle
        self._fin = weakref.finalize(self, lib.psd_sponge_free, _handle)

    @classmethod
    def new(cls, parameters: PoseidonConfig) -> "PoseidonSponge":
        return cls(parameters)

    def clone(self) -> "PoseidonSponge":
        h = C.c_void_p()
        check(lib.psd_sponge_clone(self._h, C.byref(h)))
        return PoseidonSponge(self.parameters, self._ctx.device, h)

    # -- absorb (mod.rs:232-254 + the in-scope Absorb impls of absorb.rs) ------------------------
    def absorb(self, input: AbsorbInput) -> None:
        """ints / sequences of ints are native field elements (absorb.rs:152-165, 256-280; values are
        taken mod p like F::from); bytes are a `&[u8]` (absorb.rs:122-140)."""
        if isinstance(input, (bytes, bytearray)):
            buf = (C.c_uint8 * len(input)).from_buffer_copy(bytes(input)) if len(input) else None
            check(lib.psd_sponge_absorb_bytes(self._h, buf, len(input)))
            return
        if isinstance(input, int) and not isinstance(input, bool):
            elems = [input]
        elif isinstance(input, (list, tuple)) and all(isinstance(e, int) and not isinstance(e, bool) for e in input):
            elems = list(input)
        else:   # typed stand-ins for other Rust values (sponge_b200.absorb): u8/ints/bool/Option/lists
            from .absorb import to_sponge_field_elements
            elems = to_sponge_field_elements(input, self.parameters.field)
        if not elems:
            return   # mod.rs:234-236
        p = self.parameters.modulus
        arr = ints_to_limbs(int(e) % p for e in elems)
        check(lib.psd_sponge_absorb(self._h, _ptr(arr), len(elems), L.REPR_CANONICAL))

    # -- squeezes -------------------------------------------------------------------------------
    def squeeze_native_field_elements(self, num_elements: int) -> List[int]:
        out = np.zeros((max(num_elements, 1), 4), dtype=np.uint64)
        check(lib.psd_sponge_squeeze_native(self._h, num_elements, _ptr(out), L.REPR_CANONICAL))
        return limbs_to_ints(out[:num_elements])

    def squeeze_bytes(self, num_bytes: int) -> bytes:
        out = np.zeros(max(num_bytes, 1), dtype=np.uint8)
        check(lib.psd_sponge_squeeze_bytes(self._h, num_bytes, _ptr(out)))
        return out[:num_bytes].tobytes()

    def squeeze_bits(self, num_bits: int) -> List[bool]:
        out = np.zeros(max(num_bits, 1), dtype=np.uint8)
        check(lib.psd_sponge_squeeze_bits(self._h, num_bits, _ptr(out)))
        return [bool(b) for b in out[:num_bits]]

    def squeeze_native_field_elements_with_sizes(self, sizes: Sequence[FieldElementSize]) -> List[int]:
        """lib.rs:168-182."""
        if all(s.truncated is None for s in sizes):
            return self.squeeze_native_field_elements(len(sizes))
        return self._squeeze_with_sizes_default(sizes, self.parameters.field)

    def squeeze_field_elements_with_sizes(self, sizes: Sequence[FieldElementSize], field: Optional[str] = None) -> List[int]:
        """mod.rs:288-304."""
        field = field or self.parameters.field
        if MODULUS[field] == self.parameters.modulus:
            return self.squeeze_native_field_elements_with_sizes(sizes)
        return self._squeeze_with_sizes_default(sizes, field)

    def squeeze_field_elements(self, num_elements: int, field: Optional[str] = None) -> List[int]:
        """mod.rs:306-317; `field` plays the role of the type parameter F2."""
        field = field or self.parameters.field
        if field == self.parameters.field:
            return self.squeeze_native_field_elements(num_elements)
        out = np.zeros((max(num_elements, 1), 4), dtype=np.uint64)
        check(lib.psd_sponge_squeeze_field_elements(self._h, FIELDS[field], num_elements, _ptr(out)))
        return limbs_to_ints(out[:num_elements])

    def _squeeze_with_sizes_default(self, sizes: Sequence[FieldElementSize], field: str) -> List[int]:
        """squeeze_field_elements_with_sizes_default_impl, lib.rs:61-100."""
        if len(sizes) == 0:
            return []
        widths = [s.num_bits(field) for s in sizes]
        bits = self.squeeze_bits(sum(widths))
        out, pos = [], 0
        for w in widths:
            v = 0
            for i, b in enumerate(bits[pos:pos + w]):
                v |= int(b) << i
            pos += w
            out.append(v % MODULUS[field])
        return out

    def fork(self, domain: bytes) -> "PoseidonSponge":
        """lib.rs:149-157."""
        h = C.c_void_p()
        buf = (C.c_uint8 * len(domain)).from_buffer_copy(bytes(domain)) if len(domain) else None
        check(lib.psd_sponge_fork(self._h, buf, len(domain), C.byref(h)))
        return PoseidonSponge(self.parameters, self._ctx.device, h)

    # -- SpongeExt (lib.rs:188-195, mod.rs:351-367) -------------------------------------------------
    @property
    def state(self) -> List[int]:
        return self.into_state().state

    @property
    def mode(self) -> DuplexSpongeMode:
        return self.into_state().mode

    def into_state(self) -> PoseidonSpongeState:
        st = np.zeros((self._ctx.t, 4), dtype=np.uint64)
        mode, idx = C.c_int(), C.c_uint32()
        check(lib.psd_sponge_get_state(self._h, _ptr(st), L.REPR_CANONICAL, C.byref(mode), C.byref(idx)))
        return PoseidonSpongeState(limbs_to_ints(st),
                                   DuplexSpongeMode("Absorbing" if mode.value == 0 else "Squeezing", idx.value))

    @classmethod
    def from_state(cls, state: PoseidonSpongeState, params: PoseidonConfig) -> "PoseidonSponge":
        s = cls(params)
        arr = ints_to_limbs(state.state)
        check(lib.psd_sponge_set_state(s._h, _ptr(arr), L.REPR_CANONICAL,
                                       0 if state.mode.kind == "Absorbing" else 1, state.mode.index))
        return s
