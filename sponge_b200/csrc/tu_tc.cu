// Translation units of the variant-6 kernels (integer S-box + tensor-core linear layer, poseidon_tc.cuh), one per
// field: compile with -DPSD_TU_FIELD={0,1} -DPSD_POOL_SHARED=1.  The kernels take their few integer-side
// constants (round-0 constants, R^2) from a small image staged into shared memory, next to the per-warp scratch
// tiles, and their tables from a global-memory image.
#include "registry.hpp"
#include "sponge_kernel.cuh"

#if !defined(PSD_TU_FIELD) || !PSD_POOL_KIND
#error "compile with -DPSD_TU_FIELD=<0|1> -DPSD_POOL_SHARED=1"
#endif

namespace psd {

#if PSD_TU_FIELD == 0
typedef Bn254Fr TuField;
#define PSD_TU_FNAME "bn254"
#else
typedef Bls12381Fr TuField;
#define PSD_TU_FNAME "bls12_381"
#endif

#define PSD_REG_TC(T, ALPHA)                                                                  \
    static SpongeKernelRegistrar g_reg_tc_##T##_##ALPHA(SpongeKernelEntry{                    \
        TuField::kFieldId, T, ALPHA, 6, 1, 0, PSD_MAX_BLOCK(6), sponge_kernel<TuField, T, ALPHA, 6>, \
        "sponge_kernel<" PSD_TU_FNAME ",t=" #T ",alpha=" #ALPHA ",variant=6,smem>"});

PSD_REG_TC(3, 5)
PSD_REG_TC(3, 17)
PSD_REG_TC(2, 5)
PSD_REG_TC(5, 5)
PSD_REG_TC(4, 5)

}  // namespace psd
