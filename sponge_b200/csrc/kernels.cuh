// Kernel-side argument blocks shared between kernels.cu (device) and api.cu (host).
#pragma once
#include <stddef.h>
#include <stdint.h>

#include "poseidon_core.cuh"
#include "poseidon_tc.cuh"
#include "poseidon_tc5.cuh"

namespace psd {

enum : int { kReprCanonical = 0, kReprMontgomery = 1 };
enum : int { kLayoutAos = 0, kLayoutSoa = 1 };
enum : int { kModeAbsorbing = 0, kModeSqueezing = 1 };

// One launch = every sponge of the batch runs the same (warp-uniform) call sequence:
//   [load state] -> n_raw x permute -> absorb(n_absorb) -> squeeze(n_squeeze) -> [store state]
// with the reference's bookkeeping (mod.rs:121-182, 232-254, 320-342) executed in-kernel.
struct SpongeArgs {
    PermDesc d;
    uint64_t n;             // sponges
    // state (n * T elements each).  state_in null => fresh all-zero state (PoseidonSponge::new);
    // state_out null => the final state is dropped.  The two may alias (in-place) or differ in
    // layout/representation (state export/import).
    const uint32_t* state_in;
    int32_t state_in_layout, state_in_repr;
    uint32_t* state_out;
    int32_t state_out_layout, state_out_repr;
    uint32_t n_raw;         // plain permutations applied first (psd_permute_batch)
    // absorb: n * n_absorb elements
    const uint32_t* in;
    uint32_t n_absorb;
    int32_t in_layout, in_repr;
    int32_t mode;           // DuplexSpongeMode on entry
    uint32_t index;
    // squeeze: n * n_squeeze elements, or n * n_bytes bytes when out_bytes != null
    int32_t do_squeeze;
    uint32_t n_squeeze;
    uint32_t* out;
    int32_t out_layout, out_repr;
    uint8_t* out_bytes;
    uint32_t n_bytes;
    int32_t* err;           // device flag: set to 1 if a canonical input was >= p
    // ragged kernels only: sponge i absorbs elements [in_offsets[i], in_offsets[i+1]) - in_offset_base of `in`
    // (AoS); n_absorb is ignored
    const uint64_t* in_offsets;
    uint64_t in_offset_base;
    // kernels of the shared-memory form only: the context's constant image in global memory (16-byte
    // aligned, pool_words a multiple of 4); d's offsets are relative to it
    const uint32_t* pool_img;
    uint32_t pool_words;
    // variant 6 only: the tensor-core table image in global memory and where its parts are
    const unsigned char* tc_tab;
    TcDesc tc;
    uint32_t tc_stagger;    // cycles by which the warps that share a scheduler start apart (0 = together)
};

struct LaunchCfg {
    int grid, block;
};

// Host-callable launchers of the small kernels (misc_kernels.cu).  Return cudaError_t as int.
int launch_convert(int field, const PermDesc& d, uint32_t* elems, uint64_t n, int to_mont_dir, int32_t* err,
                   LaunchCfg cfg, void* stream);
int launch_synth(int field, const PermDesc& d, uint32_t* elems, uint64_t n, uint64_t seed, uint64_t first,
                 int repr, LaunchCfg cfg, void* stream);
int launch_imad_peak(int field, int kind, uint32_t* scratch, uint32_t iters, LaunchCfg cfg, void* stream);

}  // namespace psd
