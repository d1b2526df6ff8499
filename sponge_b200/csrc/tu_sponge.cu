// One translation unit per (field, state width): compile with -DPSD_TU_FIELD={0,1} -DPSD_TU_T={2..9}.
// Keeping each instantiation in its own object lets the build run them in parallel.
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

static int tu_upload(uint32_t word_off, const uint32_t* words, uint32_t n_words, void* stream) {
    return (int)cudaMemcpyToSymbolAsync(c_pool, words, (size_t)n_words * 4, (size_t)word_off * 4,
                                        cudaMemcpyHostToDevice, (cudaStream_t)stream);
}
static ConstPoolRegistrar g_pool(TuField::kFieldId, PSD_TU_T, tu_upload);

#define PSD_REG(ALPHA, VARIANT)                                                                       \
    static SpongeKernelRegistrar g_reg_##ALPHA##_##VARIANT(SpongeKernelEntry{                         \
        TuField::kFieldId, PSD_TU_T, ALPHA, VARIANT, sponge_kernel<TuField, PSD_TU_T, ALPHA, VARIANT>, \
        "sponge_kernel<" PSD_TU_FNAME ",t=" PSD_STR(PSD_TU_T) ",alpha=" #ALPHA ",variant=" #VARIANT ">"});

PSD_REG(0, 0)
#if PSD_TU_T <= 5
PSD_REG(0, 1)
#endif
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

}  // namespace psd
