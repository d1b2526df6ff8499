// Tables for the tensor-core linear layer (variant 6; host, one-time) -- see tc_tables.cpp / poseidon_tc.cuh.
#pragma once
#include <string>

#include "opt_constants.hpp"
#include "poseidon_tc.cuh"

namespace psd {

// Builds the variant-6 table image for one parameter set (Montgomery inputs): byte-digit tables of every
// constant that multiplies the state, pre-arranged in mma.m16n8k32 B-fragment order, and the additive
// constants as accumulator start values.  td.ok is set when the variant applies (a factorable matrix, full
// rounds on both sides of the partial rounds, width <= 5) and the numeric bounds hold.
// layout 0: every table in mma.sync.m16n8k32 B-fragment order, additive constants as accumulator start values
// (variant 6).  layout 1: tables as B operands of tcgen05.mma in the K-major no-swizzle core-matrix order
// ([k / 16][column][k % 16]); the constants that multiply an S-box output have 64 rows (the operand is the unreduced
// product) and are stacked along N by operand; additive constants as eight 32-bit limbs (variant 7).
void build_tc_tables(const HostField& f, uint32_t t, uint32_t full_rounds, uint32_t partial_rounds,
                     const std::vector<Fe>& ark_mont, const std::vector<Fe>& mds_mont, TcDesc& td,
                     std::vector<uint8_t>& img, uint32_t layout = 0);

// The numeric conditions behind variant 6 for a given field and width: column sums below 2^32 / 257, sweep
// results below 2^256 and inside the S-box's operand range.  *why names the first one that fails.
bool tc_bounds_ok(const HostField& f, uint32_t t, std::string* why, uint32_t layout = 0);

// Host model of what the device computes for ONE output value (used by tests/emu): the sweep of
// sum_j bytes(x_j) * Tab(c_j) + K' followed by one 32-bit Montgomery row.  x: n_terms values of 8 words
// (any 256-bit integers), frags / kfrag as laid out by build_tc_tables.
void tc_host_row(const HostField& f, const uint32_t* x, uint32_t n_terms, const uint8_t* const* frags, const uint8_t* kfrag,
                 uint32_t* out8);

// The same for layout 1 (variant 7).  A term is a value of 32 bytes (a state element) or 64 bytes (an unreduced product)
// and the table it meets: `table` points at a stack of `stack` constants ([16-byte K chunk][constant][column][16 bytes]),
// `block` picks the constant.  kbytes: the 32 bytes of K 2^32 mod p, or null.
struct Tc5HostTerm {
    const uint32_t* value;
    uint32_t nbytes;
    const uint8_t* table;
    uint32_t stack, block;
};
// big: the sums are combined in 64 bits (the full rounds at t = 5); otherwise they must leave room for the 32-bit pair-combine
void tc5_host_row(const HostField& f, const Tc5HostTerm* terms, uint32_t n_terms, const uint8_t* kbytes, uint32_t* out8,
                  bool big = false);

}  // namespace psd
