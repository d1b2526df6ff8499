"""CPU oracle (Python big-int) for ark-sponge's Poseidon sponge.  TEST INFRASTRUCTURE ONLY.

This file restates, with Python integers, the algorithm of the reference's Poseidon path so that
the CUDA library can be checked against it.  It is NOT part of the product: only `tests/`,
`__graft_entry__.smoke()` and `bench.py`'s cpu_baseline leg may import it.  The product
(`sponge_b200`) never does, and fails loudly when its CUDA library is missing.

Parity status: BLS12-381 Fr is pinned by the 33 known-answer values held by the reference's own
tests (tests/golden/reference_kats.json, extracted by oracle/extract_reference_vectors.py).
BN254 Fr and the raw Montgomery limb representation are "parity unpinned" by the reference (it has
no BN254 type and never inspects limbs); they run through exactly the same generic code, and BN254
is additionally anchored on the public circomlib vector poseidon([1,2]).

The field arithmetic itself lives in the un-vendored dependency `ark-ff` (Cargo.toml:22-25 of the
reference, patched to unpinned git master at Cargo.toml:36-42; 0.4.0-era API).  Its published
semantics restated here: `Fp<MontBackend<_,4>,4>` holds value*2^256 mod p in four little-endian
u64 limbs, always fully reduced; `into_bigint` gives the canonical integer; `from_bigint` rejects
integers >= p; `from_le_bytes_mod_order` reduces mod p.

Each function cites the reference file:line it follows (paths relative to /root/reference/src).
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import List, Optional, Sequence, Tuple

# --------------------------------------------------------------------------------------------
# Fields.  BLS12-381 Fr modulus: test.rs:6.  BN254 Fr is not in the reference (BASELINE.json
# names it); its modulus is the standard alt_bn128 scalar-field prime.
# --------------------------------------------------------------------------------------------
BLS12_381_FR = 52435875175126190479447740508185965837690552500527637822603658699938581184513
BN254_FR = 21888242871839275222246405745257275088548364400416034343698204186575808495617

FIELD_IDS = {"bn254": 0, "bls12_381": 1}
MODULI = {"bn254": BN254_FR, "bls12_381": BLS12_381_FR}
R_BITS = 256  # Montgomery radix of Fp<MontBackend<_,4>,4>: 4 x 64-bit limbs


def modulus_bits(p: int) -> int:
    """ark-ff `MODULUS_BIT_SIZE`."""
    return p.bit_length()


def to_mont(x: int, p: int) -> int:
    """Canonical integer -> Montgomery residue (what `Fp.0.0` holds)."""
    return (x << R_BITS) % p


def from_mont(x: int, p: int) -> int:
    """Montgomery residue -> canonical integer (`into_bigint`)."""
    return (x * pow(1 << R_BITS, -1, p)) % p


def int_to_limbs64(x: int) -> List[int]:
    return [(x >> (64 * i)) & 0xFFFFFFFFFFFFFFFF for i in range(4)]


def limbs64_to_int(l: Sequence[int]) -> int:
    return sum(int(v) << (64 * i) for i, v in enumerate(l))


# --------------------------------------------------------------------------------------------
# Grain LFSR (poseidon/grain_lfsr.rs)
# --------------------------------------------------------------------------------------------
class GrainLFSR:
    """poseidon/grain_lfsr.rs:6-189."""

    def __init__(self, is_sbox_an_inverse: bool, prime_num_bits: int, state_len: int,
                 num_full_rounds: int, num_partial_rounds: int):
        # grain_lfsr.rs:15-84: 80-bit header, big-endian sub-fields
        st = [False] * 80
        st[1] = True                              # b0,b1 = 0,1: prime field
        st[5] = bool(is_sbox_an_inverse)          # b2..b5: S-box kind

        def put(lo: int, hi: int, v: int) -> None:
            for i in range(hi, lo - 1, -1):
                st[i] = (v & 1) == 1
                v >>= 1

        put(6, 17, prime_num_bits)
        put(18, 29, state_len)
        put(30, 39, num_full_rounds)
        put(40, 49, num_partial_rounds)
        for i in range(50, 80):
            st[i] = True
        self.prime_num_bits = prime_num_bits
        self.state = st
        self.head = 0
        for _ in range(160):                      # grain_lfsr.rs:176-188 (init)
            self.update()

    def update(self) -> bool:
        """grain_lfsr.rs:162-174."""
        s, h = self.state, self.head
        b = (s[(h + 62) % 80] ^ s[(h + 51) % 80] ^ s[(h + 38) % 80] ^ s[(h + 23) % 80]
             ^ s[(h + 13) % 80] ^ s[h])
        s[h] = b
        self.head = (h + 1) % 80
        return b

    def get_bits(self, n: int) -> List[bool]:
        """grain_lfsr.rs:86-106: self-shrinking pairs."""
        out = []
        for _ in range(n):
            first = self.update()
            while not first:
                self.update()
                first = self.update()
            out.append(self.update())
        return out

    def _draw_int(self) -> int:
        # bits come out most-significant first (grain_lfsr.rs:119-123, 141-155)
        v = 0
        for b in self.get_bits(self.prime_num_bits):
            v = (v << 1) | int(b)
        return v

    def field_elements_rejection_sampling(self, n: int, p: int) -> List[int]:
        """grain_lfsr.rs:108-133 (`from_bigint` -> None when >= p)."""
        assert modulus_bits(p) == self.prime_num_bits
        out = []
        for _ in range(n):
            while True:
                v = self._draw_int()
                if v < p:
                    out.append(v)
                    break
        return out

    def field_elements_mod_p(self, n: int, p: int) -> List[int]:
        """grain_lfsr.rs:135-159 (`from_le_bytes_mod_order`)."""
        assert modulus_bits(p) == self.prime_num_bits
        return [self._draw_int() % p for _ in range(n)]


# --------------------------------------------------------------------------------------------
# Parameters (poseidon/mod.rs:23-42,185-214; poseidon/traits.rs)
# --------------------------------------------------------------------------------------------
@dataclass
class PoseidonConfig:
    """poseidon/mod.rs:23-42.  Values are canonical integers in [0,p)."""
    full_rounds: int
    partial_rounds: int
    alpha: int
    mds: List[List[int]]
    ark: List[List[int]]
    rate: int
    capacity: int
    p: int = BLS12_381_FR

    def __post_init__(self):
        # shape asserts of PoseidonConfig::new, poseidon/mod.rs:196-203
        t = self.rate + self.capacity
        assert len(self.ark) == self.full_rounds + self.partial_rounds
        assert all(len(r) == t for r in self.ark)
        assert len(self.mds) == t and all(len(r) == t for r in self.mds)


def find_poseidon_ark_and_mds(p: int, prime_bits: int, rate: int, full_rounds: int,
                              partial_rounds: int, skip_matrices: int
                              ) -> Tuple[List[List[int]], List[List[int]]]:
    """poseidon/traits.rs:105-146."""
    t = rate + 1
    lfsr = GrainLFSR(False, prime_bits, t, full_rounds, partial_rounds)
    ark = [lfsr.field_elements_rejection_sampling(t, p) for _ in range(full_rounds + partial_rounds)]
    for _ in range(skip_matrices):
        lfsr.field_elements_mod_p(2 * t, p)
    xs = lfsr.field_elements_mod_p(t, p)
    ys = lfsr.field_elements_mod_p(t, p)
    mds = [[pow((xs[i] + ys[j]) % p, -1, p) for j in range(t)] for i in range(t)]
    return ark, mds


# test.rs:13-32 -- (rate, alpha, full_rounds, partial_rounds, skip_matrices)
BLS12_381_PARAMS_OPT_FOR_CONSTRAINTS = [
    (2, 17, 8, 31, 0), (3, 5, 8, 56, 0), (4, 5, 8, 56, 0), (5, 5, 8, 57, 0),
    (6, 5, 8, 57, 0), (7, 5, 8, 57, 0), (8, 5, 8, 57, 0),
]
BLS12_381_PARAMS_OPT_FOR_WEIGHTS = [(r, 257, 8, 13, 0) for r in range(2, 9)]


def get_default_poseidon_parameters(rate: int, optimized_for_weights: bool,
                                    p: int = BLS12_381_FR) -> Optional[PoseidonConfig]:
    """poseidon/traits.rs:69-102 with the BLS12-381 Fr tables of test.rs:13-32."""
    table = BLS12_381_PARAMS_OPT_FOR_WEIGHTS if optimized_for_weights \
        else BLS12_381_PARAMS_OPT_FOR_CONSTRAINTS
    for (r, alpha, rf, rp, skip) in table:
        if r == rate:
            ark, mds = find_poseidon_ark_and_mds(p, modulus_bits(p), rate, rf, rp, skip)
            return PoseidonConfig(rf, rp, alpha, mds, ark, rate, 1, p)   # capacity: 1, traits.rs:96
    return None


def bn254_config(rate: int, full_rounds: int, partial_rounds: int, alpha: int = 5) -> PoseidonConfig:
    """BASELINE.json configs C2/C3: BN254 Fr parameters through the same generator."""
    ark, mds = find_poseidon_ark_and_mds(BN254_FR, 254, rate, full_rounds, partial_rounds, 0)
    return PoseidonConfig(full_rounds, partial_rounds, alpha, mds, ark, rate, 1, BN254_FR)


# --------------------------------------------------------------------------------------------
# The sponge (poseidon/mod.rs:44-367)
# --------------------------------------------------------------------------------------------
ABSORBING = 0
SQUEEZING = 1


def permute(cfg: PoseidonConfig, state: Sequence[int]) -> List[int]:
    """poseidon/mod.rs:95-118 with apply_ark 76-80, apply_s_box 63-74, apply_mds 82-93."""
    p = cfg.p
    st = list(state)
    t = len(st)
    half = cfg.full_rounds // 2
    total = cfg.full_rounds + cfg.partial_rounds
    for r in range(total):
        full = r < half or r >= half + cfg.partial_rounds
        st = [(st[i] + cfg.ark[r][i]) % p for i in range(t)]
        if full:
            st = [pow(x, cfg.alpha, p) for x in st]
        else:
            st[0] = pow(st[0], cfg.alpha, p)
        st = [sum(st[j] * cfg.mds[i][j] for j in range(t)) % p for i in range(t)]
    return st


@dataclass
class PoseidonSponge:
    """poseidon/mod.rs:51-60; state holds canonical integers, capacity lanes first."""
    cfg: PoseidonConfig
    state: List[int] = field(default_factory=list)
    mode: int = ABSORBING
    index: int = 0
    n_permutes: int = 0   # bookkeeping for tests only

    def __post_init__(self):
        if not self.state:
            self.state = [0] * (self.cfg.rate + self.cfg.capacity)   # mod.rs:219-230

    def clone(self) -> "PoseidonSponge":
        return PoseidonSponge(self.cfg, list(self.state), self.mode, self.index, self.n_permutes)

    def _permute(self) -> None:
        self.state = permute(self.cfg, self.state)
        self.n_permutes += 1

    def _absorb_internal(self, idx: int, elems: Sequence[int]) -> None:
        """mod.rs:121-150."""
        rate, cap, p = self.cfg.rate, self.cfg.capacity, self.cfg.p
        rem = list(elems)
        while True:
            if idx + len(rem) <= rate:
                for i, e in enumerate(rem):
                    self.state[cap + i + idx] = (self.state[cap + i + idx] + e) % p
                self.mode, self.index = ABSORBING, idx + len(rem)
                return
            n = rate - idx
            for i, e in enumerate(rem[:n]):
                self.state[cap + i + idx] = (self.state[cap + i + idx] + e) % p
            self._permute()
            rem = rem[n:]
            idx = 0

    def absorb(self, elems: Sequence[int]) -> None:
        """mod.rs:232-254 for inputs that are already native field elements
        (absorb.rs:152-165, 256-280: native elements pass through unchanged)."""
        elems = [e % self.cfg.p for e in elems]
        if not elems:
            return
        if self.mode == ABSORBING:
            idx = self.index
            if idx == self.cfg.rate:
                self._permute()
                idx = 0
            self._absorb_internal(idx, elems)
        else:
            self._permute()
            self._absorb_internal(0, elems)

    def _squeeze_internal(self, idx: int, n: int) -> List[int]:
        """mod.rs:153-182, including the un-decremented length test at 175 (quirk Q1)."""
        rate, cap = self.cfg.rate, self.cfg.capacity
        out: List[int] = []
        remaining = n
        while True:
            if idx + remaining <= rate:
                out += self.state[cap + idx: cap + idx + remaining]
                self.mode, self.index = SQUEEZING, idx + remaining
                return out
            k = rate - idx
            out += self.state[cap + idx: cap + idx + k]
            if remaining != rate:
                self._permute()
            remaining -= k
            idx = 0

    def squeeze_native_field_elements(self, n: int) -> List[int]:
        """mod.rs:320-342."""
        if self.mode == ABSORBING:
            self._permute()
            return self._squeeze_internal(0, n)
        idx = self.index
        if idx == self.cfg.rate:
            self._permute()
            idx = 0
        return self._squeeze_internal(idx, n)

    def squeeze_bytes(self, num_bytes: int) -> bytes:
        """mod.rs:256-270."""
        usable = (modulus_bits(self.cfg.p) - 1) // 8
        n = (num_bytes + usable - 1) // usable
        out = b"".join(e.to_bytes(32, "little")[:usable] for e in self.squeeze_native_field_elements(n))
        return out[:num_bytes]

    def squeeze_bits(self, num_bits: int) -> List[bool]:
        """mod.rs:272-286."""
        usable = modulus_bits(self.cfg.p) - 1
        n = (num_bits + usable - 1) // usable
        bits: List[bool] = []
        for e in self.squeeze_native_field_elements(n):
            bits += [bool((e >> i) & 1) for i in range(usable)]
        return bits[:num_bits]

    def squeeze_field_elements(self, n: int, p2: Optional[int] = None) -> List[int]:
        """mod.rs:288-317 + lib.rs:61-100, 168-182.  `p2` = the target field's modulus.
        FieldElementSize::num_bits always returns MODULUS_BIT_SIZE-1 (lib.rs:45-52, quirk Q3),
        so sizes do not matter and are not modelled."""
        if p2 is None or p2 == self.cfg.p:
            return self.squeeze_native_field_elements(n)
        if n == 0:
            return []
        nb = modulus_bits(p2) - 1
        bits = self.squeeze_bits(nb * n)
        out = []
        for k in range(n):
            v = 0
            for i, b in enumerate(bits[k * nb:(k + 1) * nb]):
                v |= int(b) << i
            out.append(v % p2)
        return out

    # -- absorb.rs marshalling for non-field inputs ("next" row f1) -----------------------------
    def absorb_bytes(self, data: bytes) -> None:
        """absorb.rs:122-140: `[u8]`/`Vec<u8>` -> u64-LE length prefix, then 31-byte LE chunks
        (ark-ff `ToConstraintField<F> for [u8]`, un-vendored: chunk = (MODULUS_BIT_SIZE-1)/8 bytes,
        each chunk read with from_le_bytes_mod_order)."""
        self.absorb(bytes_to_field_elements(data, self.cfg.p))

    def fork(self, domain: bytes) -> "PoseidonSponge":
        """lib.rs:149-157: clone, then absorb (usize len as u64 LE bytes || domain) as Vec<u8>."""
        s = self.clone()
        s.absorb_bytes(len(domain).to_bytes(8, "little") + domain)
        return s


def bytes_to_field_elements(data: bytes, p: int) -> List[int]:
    """absorb.rs:135-139."""
    usable = (modulus_bits(p) - 1) // 8
    buf = len(data).to_bytes(8, "little") + bytes(data)
    return [int.from_bytes(buf[i:i + usable], "little") % p for i in range(0, len(buf), usable)]


# --------------------------------------------------------------------------------------------
# Batch shapes added by the build -- defined *via* the reference sponge.
# --------------------------------------------------------------------------------------------
def hash_absorb_squeeze(cfg: PoseidonConfig, inputs: Sequence[int], n_out: int) -> List[int]:
    s = PoseidonSponge(cfg)
    s.absorb(inputs)
    return s.squeeze_native_field_elements(n_out)


def compress2(cfg: PoseidonConfig, l: int, r: int) -> int:
    """H(l,r) := new; absorb([l,r]); squeeze_native(1)[0]  (rate >= 2)."""
    return hash_absorb_squeeze(cfg, [l, r], 1)[0]


def merkle_root(cfg: PoseidonConfig, leaves: Sequence[int]) -> int:
    level = list(leaves)
    assert len(level) >= 1 and (len(level) & (len(level) - 1)) == 0
    while len(level) > 1:
        level = [compress2(cfg, level[i], level[i + 1]) for i in range(0, len(level), 2)]
    return level[0]


# --------------------------------------------------------------------------------------------
# Synthetic inputs shared by tests and bench: counter-based splitmix64.
# --------------------------------------------------------------------------------------------
_M64 = 0xFFFFFFFFFFFFFFFF


def splitmix64(x: int) -> int:
    x = (x + 0x9E3779B97F4A7C15) & _M64
    z = x
    z = ((z ^ (z >> 30)) * 0xBF58476D1CE4E5B9) & _M64
    z = ((z ^ (z >> 27)) * 0x94D049BB133111EB) & _M64
    return z ^ (z >> 31)


def synth_element(seed: int, k: int, p: int) -> int:
    """Element k of stream `seed`: 4 splitmix64 words -> 256-bit LE, top bits cleared to
    bits(p), redrawn with a salt until < p.  Canonical, uniform in [0,p)."""
    nbits = modulus_bits(p)
    salt = 0
    while True:
        v = 0
        for i in range(4):
            ctr = ((seed << 40) ^ (salt << 34)) + 4 * k + i
            v |= splitmix64(ctr & _M64) << (64 * i)
        v &= (1 << nbits) - 1
        if v < p:
            return v
        salt += 1
