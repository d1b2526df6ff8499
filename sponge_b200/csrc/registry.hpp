// Every translation unit that holds kernels has its own copy of the constant bank image (c_pool is
// a per-module __constant__ symbol) and registers (a) an uploader for it and (b) the sponge kernels
// it instantiates.  api.cu walks these tables; nothing here is device code.
#pragma once
#include <stdint.h>

#include "kernels.cuh"

namespace psd {

typedef int (*PoolUploadFn)(uint32_t word_off, const uint32_t* words, uint32_t n_words, void* stream);
typedef void (*SpongeKernelFn)(const SpongeArgs);

struct SpongeKernelEntry {
    int field, t;
    uint64_t alpha;     // 0 = any alpha (runtime ladder)
    int variant;        // 0 = dense, 1 = optimised
    SpongeKernelFn fn;
    const char* name;
};

void register_pool_uploader(PoolUploadFn fn);
void register_sponge_kernel(const SpongeKernelEntry& e);

struct ConstPoolRegistrar {
    explicit ConstPoolRegistrar(PoolUploadFn fn) { register_pool_uploader(fn); }
};
struct SpongeKernelRegistrar {
    explicit SpongeKernelRegistrar(const SpongeKernelEntry& e) { register_sponge_kernel(e); }
};

}  // namespace psd
