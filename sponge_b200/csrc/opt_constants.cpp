// Host-side derivation of constants for the optimised permutation kernels.
#include "opt_constants.hpp"

namespace psd {

void append_optimised_constants(const HostField& f, uint32_t t, uint32_t full_rounds, uint32_t partial_rounds,
                                const std::vector<Fe>& ark_mont, const std::vector<Fe>& mds_mont, PermDesc& d,
                                std::vector<uint32_t>& img) {
    (void)f; (void)t; (void)full_rounds; (void)partial_rounds; (void)ark_mont; (void)mds_mont; (void)d; (void)img;
}

void offset_desc(PermDesc& d, uint32_t pool_off) {
    d.ark_off += pool_off;
    d.mds_off += pool_off;
    d.one_off += pool_off;
    d.r2_off += pool_off;
}

}  // namespace psd
