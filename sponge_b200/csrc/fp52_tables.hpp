// Tables for the FP64-pipe linear layer (host, one-time) -- see fp52_tables.cpp.
#pragma once
#include <string>

#include "opt_constants.hpp"

namespace psd {

// Appends the variant-5 constants for one parameter set (Montgomery inputs) to `img` and records their
// offsets (8-byte units, image-relative) in d.fp_*; sets d.fp_ok when the variant applies (BN254 Fr, a
// factorable matrix, full rounds on both sides of the partial rounds).  Independent of d.opt_ok /
// d.ctab_ok: the image only needs ark/mds/one/r2 besides these tables.
void append_fp52_tables(const HostField& f, uint32_t t, uint32_t full_rounds, uint32_t partial_rounds,
                        const std::vector<Fe>& ark_mont, const std::vector<Fe>& mds_mont, PermDesc& d,
                        std::vector<uint32_t>& img);

// Evaluates, for the image append_fp52_tables produced (or any image with the same layout), the numeric bounds
// that make variant 5 exact -- table limbs below 2^52, entries below p, rows below p + 2^207, lazily reduced
// lanes below 2^258, no column-sum wrap, the lazy S-box range -- with the worst case of every state limb
// against the ACTUAL table entries.  append_fp52_tables clears d.fp_ok when it fails (the context then runs
// an integer variant).  *why (optional) names the first bound that failed.
bool fp52_image_bounds_ok(const HostField& f, uint32_t t, uint32_t full_rounds, uint32_t partial_rounds, const PermDesc& d,
                          const std::vector<uint32_t>& img, std::string* why);

}  // namespace psd
