// Host-only parameter generation (see params.cpp).
#pragma once
#include "host_field.hpp"

namespace psd {

struct DefaultEntry {
    uint32_t rate, full_rounds, partial_rounds, skip_matrices;
    uint64_t alpha;
};

// find_poseidon_ark_and_mds, /root/reference/src/poseidon/traits.rs:105-146.  Canonical outputs:
// ark[(rf+rp)*(rate+1)], mds[(rate+1)^2].  Returns 0, -1 (prime_bits mismatch), -2 (singular).
int find_poseidon_ark_and_mds(const HostField& f, uint64_t prime_bits, uint32_t rate, uint32_t rf,
                              uint32_t rp, uint32_t skip_matrices, Fe* ark, Fe* mds);

// Default table lookup, traits.rs:69-102 / test.rs:13-32.
bool default_entry(int field_id, uint32_t rate, bool weights, DefaultEntry* e);

}  // namespace psd
