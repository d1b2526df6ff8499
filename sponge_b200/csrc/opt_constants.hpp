// Derivation of the optimised-permutation constants (host, one-time) -- see opt_constants.cpp.
#pragma once
#include <stdint.h>

#include <vector>

#include "host_field.hpp"
#include "poseidon_core.cuh"

namespace psd {

// Appends the optimised-permutation constants for one parameter set (Montgomery inputs) to the
// constant-bank image and records their image-relative offsets in `d`.
void append_optimised_constants(const HostField& f, uint32_t t, uint32_t full_rounds, uint32_t partial_rounds,
                                const std::vector<Fe>& ark_mont, const std::vector<Fe>& mds_mont, PermDesc& d,
                                std::vector<uint32_t>& img);

// Appends the digit tables of every round constant that multiplies the state (const_dot in
// fp256.cuh) and sets d.ctab_ok.  Needs the optimised form (d.opt_ok); roughly 8x the size of the
// constants it covers, so callers may drop it again (truncate `img`, clear ctab_ok) if the constant
// bank is short.
// lanes_only: tables for the lane-update constants v_j only (sets ctab_ok = 2).
void append_digit_tables(const HostField& f, uint32_t t, uint32_t full_rounds, uint32_t partial_rounds,
                         const std::vector<Fe>& mds_mont, PermDesc& d, std::vector<uint32_t>& img, bool lanes_only = false);

// Rebases every offset of `d` from image-relative to pool-relative.
void offset_desc(PermDesc& d, uint32_t pool_off);

}  // namespace psd
