"""Host-side marshalling of non-field values into sponge input -- mirror of ark-sponge's `Absorb`
impls (reference src/absorb.rs), row f1 of SURVEY.md 8(f).  Pure conversion of Python stand-ins for
Rust values into bytes / native field elements; the sponge itself still runs on the GPU.

Rust type            stand-in here                 reference
-------------------  ----------------------------  ---------------------------------------------
u8                   U8(v)                         absorb.rs:122-140
bool                 Bool(v) / Python bool         absorb.rs:142-150
Fp (native element)  Fe(v) / Python int            absorb.rs:152-165  (via field_cast, 106-120)
u16..u128            U16(v) .. U128(v)             absorb.rs:167-184
i8..i128             I8(v) .. I128(v)              absorb.rs:186-208
usize / isize        Usize(v) / Isize(v)           absorb.rs:210-228  (as u64 / i64)
&[A], Vec<A>         Python list / tuple / bytes   absorb.rs:256-286  (A::batch_to_sponge_*)
Option<A>            Some(x) / NONE                absorb.rs:288-302
AbsorbWithLength     WithLength(list)              absorb.rs:82-101
Curve points (absorb.rs:230-254) are out of scope: they need all of ark-ec.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Any, List, Sequence

from .poseidon import MODULUS, MODULUS_BIT_SIZE


@dataclass(frozen=True)
class _Uint:
    v: int
    BITS = 0

    def __post_init__(self):
        if not 0 <= self.v < (1 << self.BITS):
            raise OverflowError(f"{self.v} does not fit u{self.BITS}")


@dataclass(frozen=True)
class _Sint:
    v: int
    BITS = 0

    def __post_init__(self):
        if not -(1 << (self.BITS - 1)) <= self.v < (1 << (self.BITS - 1)):
            raise OverflowError(f"{self.v} does not fit i{self.BITS}")


def _mk(base, name, bits):
    return type(name, (base,), {"BITS": bits, "__doc__": f"Rust {name.lower()}"})


U8, U16, U32, U64, U128 = (_mk(_Uint, n, b) for n, b in (("U8", 8), ("U16", 16), ("U32", 32), ("U64", 64), ("U128", 128)))
I8, I16, I32, I64, I128 = (_mk(_Sint, n, b) for n, b in (("I8", 8), ("I16", 16), ("I32", 32), ("I64", 64), ("I128", 128)))
Usize = _mk(_Uint, "Usize", 64)     # absorbed `as u64` (absorb.rs:210-218)
Isize = _mk(_Sint, "Isize", 64)     # absorbed `as i64` (absorb.rs:220-228)


@dataclass(frozen=True)
class Bool:
    v: bool


@dataclass(frozen=True)
class Fe:
    """A native field element (canonical integer)."""
    v: int


@dataclass(frozen=True)
class Some:
    x: Any


class _NoneType:
    def __repr__(self):
        return "NONE"


NONE = _NoneType()


@dataclass(frozen=True)
class WithLength:
    """AbsorbWithLength for a slice/Vec (absorb.rs:82-101, 265-269, 282-286)."""
    items: Sequence[Any]


def _norm(x):
    if isinstance(x, bool):
        return Bool(x)
    if isinstance(x, int):
        return Fe(x)
    if isinstance(x, (bytes, bytearray)):
        return [U8(b) for b in x]
    return x


# ---------------------------------------------------------------------------------------------
# to_sponge_bytes
# ---------------------------------------------------------------------------------------------
def to_sponge_bytes(x, field: str = "bls12_381") -> bytes:
    x = _norm(x)
    if isinstance(x, _Uint):
        return x.v.to_bytes(x.BITS // 8, "little")                      # to_le_bytes
    if isinstance(x, _Sint):
        return (x.v & ((1 << x.BITS) - 1)).to_bytes(x.BITS // 8, "little")
    if isinstance(x, Bool):
        return bytes([int(x.v)])
    if isinstance(x, Fe):
        return (x.v % MODULUS[field]).to_bytes(32, "little")            # serialize_compressed: 32-byte LE canonical
    if isinstance(x, Some):
        return b"\x01" + to_sponge_bytes(x.x, field)
    if x is NONE:
        return b"\x00"
    if isinstance(x, WithLength):
        return to_sponge_bytes(Usize(len(x.items)), field) + to_sponge_bytes(list(x.items), field)
    if isinstance(x, (list, tuple)):
        return b"".join(to_sponge_bytes(e, field) for e in x)           # batch_to_sponge_bytes
    raise TypeError(f"no Absorb stand-in for {type(x).__name__}")


# ---------------------------------------------------------------------------------------------
# to_sponge_field_elements
# ---------------------------------------------------------------------------------------------
def bytes_to_field_elements(data: bytes, field: str) -> List[int]:
    """u8::batch_to_sponge_field_elements (absorb.rs:135-139): u64-LE length prefix, then chunks of
    (MODULUS_BIT_SIZE-1)/8 bytes, each a little-endian integer (ark-ff ToConstraintField<[u8]>)."""
    usable = (MODULUS_BIT_SIZE[field] - 1) // 8
    buf = len(data).to_bytes(8, "little") + bytes(data)
    return [int.from_bytes(buf[i:i + usable], "little") % MODULUS[field] for i in range(0, len(buf), usable)]


def to_sponge_field_elements(x, field: str = "bls12_381") -> List[int]:
    p = MODULUS[field]
    if isinstance(x, (bytes, bytearray)):
        # a u8 slice even when empty: the length prefix alone still yields one element (absorb.rs:135-139)
        return bytes_to_field_elements(bytes(x), field)
    x = _norm(x)
    if isinstance(x, _Uint):
        return [x.v % p]                                                 # F::from(x)
    if isinstance(x, _Sint):
        return [abs(x.v) % p if x.v >= 0 else (-(abs(x.v) % p)) % p]      # F::from(|x|), negated if x < 0
    if isinstance(x, Bool):
        return [int(x.v)]
    if isinstance(x, Fe):
        return [x.v % p]
    if isinstance(x, Some):
        return [1] + to_sponge_field_elements(x.x, field)
    if x is NONE:
        return [0]
    if isinstance(x, WithLength):
        return to_sponge_field_elements(Usize(len(x.items)), field) + to_sponge_field_elements(list(x.items), field)
    if isinstance(x, (list, tuple)):
        items = [_norm(e) for e in x]
        # (an EMPTY Python list carries no element type; it is treated as a non-u8 slice -> no elements.
        #  Pass b"" for an empty byte slice.)
        if items and all(isinstance(e, _Uint) and e.BITS == 8 for e in items):
            return bytes_to_field_elements(bytes(e.v for e in items), field)   # the u8 batch rule
        out: List[int] = []
        for e in items:
            out += to_sponge_field_elements(e, field)
        return out
    raise TypeError(f"no Absorb stand-in for {type(x).__name__}")


# ---------------------------------------------------------------------------------------------
# macros (absorb.rs:319-354)
# ---------------------------------------------------------------------------------------------
def absorb_all(sponge, *absorbables) -> None:
    """absorb!(sponge, a_0, ..., a_n): one absorb call per argument."""
    for a in absorbables:
        sponge.absorb(to_sponge_field_elements(a, sponge.parameters.field))


def collect_sponge_bytes(*absorbables, field: str = "bls12_381") -> bytes:
    return b"".join(to_sponge_bytes(a, field) for a in absorbables)


def collect_sponge_field_elements(*absorbables, field: str = "bls12_381") -> List[int]:
    out: List[int] = []
    for a in absorbables:
        out += to_sponge_field_elements(a, field)
    return out
