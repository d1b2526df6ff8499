"""Marshalling of non-field values (reference src/absorb.rs) -- host-side, no GPU.  Values follow
from the reference's definitions; the sponge-level behaviour (tests.rs:26-117) is in the gpu tests."""
import pytest


@pytest.fixture(scope="module")
def A():
    from sponge_b200 import absorb
    return absorb


P = 52435875175126190479447740508185965837690552500527637822603658699938581184513


def test_scalars(A):
    assert A.to_sponge_field_elements(A.U8(7)) == [7]
    assert A.to_sponge_field_elements(A.U64(2**64 - 1)) == [2**64 - 1]
    assert A.to_sponge_field_elements(A.U128(2**128 - 1)) == [2**128 - 1]
    assert A.to_sponge_field_elements(A.I16(-5)) == [P - 5]            # absorb.rs:193-199
    assert A.to_sponge_field_elements(A.I64(-(2**63))) == [P - 2**63]
    assert A.to_sponge_field_elements(A.Isize(-1)) == [P - 1]
    assert A.to_sponge_field_elements(A.Usize(3)) == [3]
    assert A.to_sponge_field_elements(True) == [1] and A.to_sponge_field_elements(A.Bool(False)) == [0]
    assert A.to_sponge_field_elements(12345) == [12345]                # native element passes through
    assert A.to_sponge_bytes(A.U32(0x01020304)) == bytes([4, 3, 2, 1])
    assert A.to_sponge_bytes(A.I16(-2)) == bytes([0xfe, 0xff])
    assert A.to_sponge_bytes(A.Usize(5)) == (5).to_bytes(8, "little")
    assert A.to_sponge_bytes(A.Fe(1)) == (1).to_bytes(32, "little")    # serialize_compressed
    with pytest.raises(OverflowError):
        A.U8(256)


def test_byte_slices_get_length_prefix_and_31_byte_chunks(A):
    """absorb.rs:135-139"""
    data = bytes(range(100))
    buf = (100).to_bytes(8, "little") + data
    want = [int.from_bytes(buf[i:i + 31], "little") for i in range(0, 108, 31)]
    assert A.to_sponge_field_elements(data) == want and len(want) == 4
    assert A.to_sponge_field_elements([A.U8(b) for b in data]) == want
    assert A.to_sponge_bytes(data) == data                              # batch_to_sponge_bytes: no prefix
    assert A.to_sponge_field_elements(b"", "bn254") == [0]              # empty u8 slice: the length prefix alone
    assert A.to_sponge_field_elements([], "bn254") == []                # untyped empty list: nothing
    # BN254: same 31-byte chunks (253 usable bits)
    assert A.to_sponge_field_elements(data, "bn254") == want


def test_option_list_and_with_length(A):
    assert A.to_sponge_field_elements(A.NONE) == [0]
    assert A.to_sponge_field_elements(A.Some(A.U16(9))) == [1, 9]
    assert A.to_sponge_bytes(A.Some(A.U8(9))) == bytes([1, 9]) and A.to_sponge_bytes(A.NONE) == b"\x00"
    assert A.to_sponge_field_elements([A.I32(1), A.I32(-1)]) == [1, P - 1]
    # list_with_nonconstant_size_element (tests.rs:45-69): the two nestings must differ
    l1 = [A.WithLength([A.U8(v) for v in (1, 2, 3, 4)]), A.WithLength([A.U8(5), A.U8(6)])]
    l2 = [A.WithLength([A.U8(1), A.U8(2)]), A.WithLength([A.U8(v) for v in (3, 4, 5, 6)])]
    assert A.to_sponge_bytes(l1) != A.to_sponge_bytes(l2)
    assert A.to_sponge_field_elements(l1) != A.to_sponge_field_elements(l2)
    assert A.to_sponge_field_elements(l1)[0] == 4                       # absorb_length first


def test_macros(A):
    """tests.rs:104-116: collect_sponge_* == manual concatenation."""
    v = [A.I32(x) for x in (6, 5, 4, 3, 2, 1)]
    assert A.collect_sponge_bytes(v, A.Fe(42)) == A.to_sponge_bytes(v) + A.to_sponge_bytes(A.Fe(42))
    assert A.collect_sponge_field_elements(v, A.Fe(42)) == [6, 5, 4, 3, 2, 1, 42]
