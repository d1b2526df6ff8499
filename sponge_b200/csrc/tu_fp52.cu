// Translation units of the variant-5 kernels (integer S-box + FP64-pipe linear layer, poseidon_fp52.cuh),
// one per field: compile with -DPSD_TU_FIELD={0,1}.  Each has its own constant pool: the 52-bit-limb
// tables of one parameter set take about 54 KiB, which would not fit next to the integer kernels'
// constants and digit tables of tu_sponge.cu.  Registered under the pool key (field, kFp52PoolKey);
// see api.cu.  The shipped library holds the BN254 unit only: for BLS12-381 (alpha = 17 S-boxes dominate,
// the S-box cannot be lazy, and the integer path already has the cheap 6-multiply Montgomery step of
// that modulus) this variant passed the parity tests but ran c5 at 140 M perm/s against 168 M for variant 3.
#include "registry.hpp"
#include "sponge_kernel.cuh"

#if !defined(PSD_TU_FIELD)
#error "compile with -DPSD_TU_FIELD=<0|1>"
#endif

namespace psd {

#if PSD_TU_FIELD == 0
typedef Bn254Fr TuField;
#define PSD_TU_FNAME "bn254"
#else
typedef Bls12381Fr TuField;
#define PSD_TU_FNAME "bls12_381"
#endif

#if PSD_POOL_KIND == 0
static int tu_upload(uint32_t word_off, const uint32_t* words, uint32_t n_words, void* stream) {
    return (int)cudaMemcpyToSymbolAsync(c_pool, words, (size_t)n_words * 4, (size_t)word_off * 4,
                                        cudaMemcpyHostToDevice, (cudaStream_t)stream);
}

// one pool for all widths of this unit: a parameter set fills most of it anyway
static ConstPoolRegistrar g_pool(TuField::kFieldId, kFp52PoolKey, tu_upload);
#define PSD_POOL_TAG ""
#else
#define PSD_POOL_TAG ",smem"
#endif

#define PSD_REG_FP52(T, ALPHA)                                                                 \
    static SpongeKernelRegistrar g_reg_##T##_##ALPHA(SpongeKernelEntry{                        \
        TuField::kFieldId, T, ALPHA, 5, PSD_POOL_KIND, 0, PSD_MAX_BLOCK(5), sponge_kernel<TuField, T, ALPHA, 5>, \
        "sponge_kernel<" PSD_TU_FNAME ",t=" #T ",alpha=" #ALPHA ",variant=5" PSD_POOL_TAG ">"}); \
    static SpongeKernelRegistrar g_regr_##T##_##ALPHA(SpongeKernelEntry{                       \
        TuField::kFieldId, T, ALPHA, 5, PSD_POOL_KIND, 1, PSD_MAX_BLOCK(5),                    \
        sponge_kernel<TuField, T, ALPHA, 5, PSD_POOL_KIND, true>,                              \
        "sponge_kernel<" PSD_TU_FNAME ",t=" #T ",alpha=" #ALPHA ",variant=5" PSD_POOL_TAG ",ragged>"});

PSD_REG_FP52(3, 5)
PSD_REG_FP52(3, 17)
#if PSD_TU_FIELD == 0
PSD_REG_FP52(2, 5)
#endif
// t = 5: the tables (about 96 KiB) exceed the constant bank, so this width exists in the shared-memory
// form only
#if PSD_POOL_KIND == 1 && PSD_TU_FIELD == 0
PSD_REG_FP52(5, 5)
#endif

}  // namespace psd
