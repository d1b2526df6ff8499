"""sponge_b200 -- B200-native batched Poseidon sponge engine (drop-in for the Poseidon path of
arkworks-rs/sponge).  Thin ctypes host layer over libposeidon_b200.so; see include/poseidon_b200.h.
"""
from ._lib import (LAYOUT_AOS, LAYOUT_SOA, MEM_DEVICE, MEM_HOST, REPR_CANONICAL, REPR_MONTGOMERY,  # noqa: F401
                   PoseidonB200Error, LIB_PATH)
from .poseidon import (DuplexSpongeMode, FieldElementSize, PoseidonConfig, PoseidonContext,  # noqa: F401
                       PoseidonParameters, PoseidonSponge, PoseidonSpongeState, SpongeBatch,
                       find_poseidon_ark_and_mds, get_default_poseidon_parameters, ints_to_limbs,
                       MultiContext, shard_range,
                       kernel_launch_count, limbs_to_ints, MODULUS, MODULUS_BIT_SIZE)

from . import absorb  # noqa: F401,E402

__version__ = "0.1.0"
