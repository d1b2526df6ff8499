// Translation units of the variant-7 kernels (integer S-box + linear layer through tcgen05.mma / TMEM, poseidon_tc5.cuh),
// one per field: compile with -DPSD_TU_FIELD={0,1} -DPSD_POOL_SHARED=1.  Like variant 6 the kernels stage only the round-0
// constants and R^2 from a small image; the tables come from a global-memory image in the tcgen05 B-operand layout.
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

#define PSD_REG_TC5(T, ALPHA)                                                                 \
    static SpongeKernelRegistrar g_reg_tc5_##T##_##ALPHA(SpongeKernelEntry{                   \
        TuField::kFieldId, T, ALPHA, 7, 1, 0, PSD_MAX_BLOCK_T(T, 7), sponge_kernel<TuField, T, ALPHA, 7>, \
        "sponge_kernel<" PSD_TU_FNAME ",t=" #T ",alpha=" #ALPHA ",variant=7,smem>"});

// the same kernels with per-sponge absorb lengths (ragged batches): the block permutes while any of its sponges needs to
#define PSD_REG_TC5_RAGGED(T, ALPHA)                                                                   \
    static SpongeKernelRegistrar g_reg_tc5r_##T##_##ALPHA(SpongeKernelEntry{                            \
        TuField::kFieldId, T, ALPHA, 7, 1, 1, PSD_MAX_BLOCK_T(T, 7), sponge_kernel<TuField, T, ALPHA, 7, 1, true>, \
        "sponge_kernel<" PSD_TU_FNAME ",t=" #T ",alpha=" #ALPHA ",variant=7,smem,ragged>"});

PSD_REG_TC5_RAGGED(3, 5)
PSD_REG_TC5_RAGGED(3, 17)
PSD_REG_TC5_RAGGED(5, 5)
PSD_REG_TC5_RAGGED(2, 5)
PSD_REG_TC5_RAGGED(4, 5)
PSD_REG_TC5(3, 5)
PSD_REG_TC5(3, 17)
PSD_REG_TC5(2, 5)
PSD_REG_TC5(4, 5)
PSD_REG_TC5(5, 5)

}  // namespace psd
