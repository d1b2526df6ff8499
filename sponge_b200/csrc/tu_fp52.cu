// Translation unit of the variant-5 kernels (integer S-box + FP64-pipe linear layer, poseidon_fp52.cuh).
// It has its own constant pool: the 52-bit-limb tables of one parameter set take about 54 KiB, which
// would not fit next to the integer kernels' constants and digit tables of tu_sponge.cu.
// Registered under the pool key (field, kFp52PoolKey); see api.cu.
#include "registry.hpp"
#include "sponge_kernel.cuh"

namespace psd {

static int tu_upload(uint32_t word_off, const uint32_t* words, uint32_t n_words, void* stream) {
    return (int)cudaMemcpyToSymbolAsync(c_pool, words, (size_t)n_words * 4, (size_t)word_off * 4,
                                        cudaMemcpyHostToDevice, (cudaStream_t)stream);
}

// one pool for all widths of this unit: a parameter set fills most of it anyway
static ConstPoolRegistrar g_pool(Bn254Fr::kFieldId, kFp52PoolKey, tu_upload);

#define PSD_REG_FP52(T, ALPHA)                                                                 \
    static SpongeKernelRegistrar g_reg_##T##_##ALPHA(SpongeKernelEntry{                        \
        Bn254Fr::kFieldId, T, ALPHA, 5, sponge_kernel<Bn254Fr, T, ALPHA, 5>,                   \
        "sponge_kernel<bn254,t=" #T ",alpha=" #ALPHA ",variant=5>"});

PSD_REG_FP52(3, 5)
PSD_REG_FP52(3, 17)
PSD_REG_FP52(2, 5)

}  // namespace psd
