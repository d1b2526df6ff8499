// One translation unit per (field, state width): compile with -DPSD_TU_FIELD={0,1} -DPSD_TU_T={2..9}.
// Keeping each instantiation in its own object lets the build run them in parallel.
// A second object per (field, width) is built with -DPSD_POOL_SHARED=1: the same kernels reading their
// constants from a per-block shared-memory copy of the context's image instead of the constant bank.
#include "registry.hpp"
#include "sponge_kernel.cuh"

#if !defined(PSD_TU_FIELD) || !defined(PSD_TU_T)
#error "compile with -DPSD_TU_FIELD=<0|1> -DPSD_TU_T=<2..9>"
#endif

namespace psd {

#if PSD_TU_FIELD == 0
typedef Bn254Fr TuField;
#define PSD_TU_FNAME "bn254"
#else
typedef Bls12381Fr TuField;
#define PSD_TU_FNAME "bls12_381"
#endif

#define PSD_STR2(x) #x
#define PSD_STR(x) PSD_STR2(x)

#if PSD_POOL_KIND == 0
static int tu_upload(uint32_t word_off, const uint32_t* words, uint32_t n_words, void* stream) {
    return (int)cudaMemcpyToSymbolAsync(c_pool, words, (size_t)n_words * 4, (size_t)word_off * 4,
                                        cudaMemcpyHostToDevice, (cudaStream_t)stream);
}
static ConstPoolRegistrar g_pool(TuField::kFieldId, PSD_TU_T, tu_upload);
#define PSD_POOL_TAG ""
#else
#define PSD_POOL_TAG ",smem"
#endif

#define PSD_REG(ALPHA, VARIANT)                                                                       \
    static SpongeKernelRegistrar g_reg_##ALPHA##_##VARIANT(SpongeKernelEntry{                         \
        TuField::kFieldId, PSD_TU_T, ALPHA, VARIANT, PSD_POOL_KIND, 0, PSD_MAX_BLOCK(VARIANT),        \
        sponge_kernel<TuField, PSD_TU_T, ALPHA, VARIANT>,                                             \
        "sponge_kernel<" PSD_TU_FNAME ",t=" PSD_STR(PSD_TU_T) ",alpha=" #ALPHA ",variant=" #VARIANT PSD_POOL_TAG ">"});
// the same kernel with per-sponge absorb lengths (ragged batches): registered for the forms a context ends
// up running by default, so that a ragged call never changes the algorithm
#define PSD_REG_RAGGED(ALPHA, VARIANT)                                                                \
    static SpongeKernelRegistrar g_regr_##ALPHA##_##VARIANT(SpongeKernelEntry{                        \
        TuField::kFieldId, PSD_TU_T, ALPHA, VARIANT, PSD_POOL_KIND, 1, PSD_MAX_BLOCK(VARIANT),        \
        sponge_kernel<TuField, PSD_TU_T, ALPHA, VARIANT, PSD_POOL_KIND, true>,                        \
        "sponge_kernel<" PSD_TU_FNAME ",t=" PSD_STR(PSD_TU_T) ",alpha=" #ALPHA ",variant=" #VARIANT PSD_POOL_TAG ",ragged>"});

PSD_REG(0, 0)
PSD_REG_RAGGED(0, 0)
#if PSD_TU_T <= 5
PSD_REG(0, 1)
PSD_REG_RAGGED(0, 1)
#endif
#if PSD_TU_T == 3
PSD_REG_RAGGED(5, 3)
PSD_REG_RAGGED(17, 3)
#endif
#if PSD_TU_T == 5
PSD_REG_RAGGED(5, 4)
#endif
#if PSD_POOL_KIND == 0
#if PSD_TU_T == 3 || PSD_TU_T == 5
PSD_REG(5, 0)
PSD_REG(5, 1)
PSD_REG(5, 2)
#endif
#if PSD_TU_T == 5
PSD_REG(5, 4)
#endif
#if PSD_TU_T == 3
PSD_REG(5, 3)
PSD_REG(17, 0)
PSD_REG(17, 1)
PSD_REG(17, 2)
PSD_REG(17, 3)
#endif
#else
// Shared-memory form: the dense kernel for every width (so that no valid parameter set is ever refused
// for lack of constant-bank space) and, for the benchmark widths, the fastest integer form with its
// full digit tables -- the image is not limited to 62 KiB here.
#if PSD_TU_T == 3
PSD_REG(5, 3)
PSD_REG(17, 3)
#endif
#if PSD_TU_T == 5
PSD_REG(5, 4)
#endif
#endif

}  // namespace psd
