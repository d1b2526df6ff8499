// Every translation unit that holds kernels has its OWN 62 KiB constant pool (c_pool is a per-unit
// __constant__ symbol) and registers (a) an uploader for it and (b) the sponge kernels it
// instantiates.  A context's constants are uploaded only into the pool of the unit whose kernels it
// runs, so parameter sets of different (field, t) never compete for constant-bank space.
// api.cu walks these tables; nothing here is device code.
#pragma once
#include <stdint.h>

#include "kernels.cuh"

namespace psd {

typedef int (*PoolUploadFn)(uint32_t word_off, const uint32_t* words, uint32_t n_words, void* stream);
typedef void (*SpongeKernelFn)(const SpongeArgs);

struct SpongeKernelEntry {
    int field, t;
    uint64_t alpha;     // 0 = any alpha (runtime ladder)
    int variant;        // 0 = dense ... 5 = FP64-pipe linear layer (see psd_ctx_set_variant)
    int pool;           // 0 = constants in the unit's constant-bank pool, 1 = staged into shared memory
    int ragged;         // 1 = per-sponge absorb lengths (SpongeArgs::in_offsets)
    int max_block;      // largest block size the kernel was compiled for (its __launch_bounds__)
    SpongeKernelFn fn;
    const char* name;
};

// Pool key of the FP64-pipe unit (tu_fp52.cu), used in place of t for its pool and uploader.
constexpr int kFp52PoolKey = 10;

// field/t identify the translation unit's kernels; the small-kernel unit registers as (-1, 0)
void register_pool_uploader(int field, int t, PoolUploadFn fn);
void register_sponge_kernel(const SpongeKernelEntry& e);

struct ConstPoolRegistrar {
    ConstPoolRegistrar(int field, int t, PoolUploadFn fn) { register_pool_uploader(field, t, fn); }
};
struct SpongeKernelRegistrar {
    explicit SpongeKernelRegistrar(const SpongeKernelEntry& e) { register_sponge_kernel(e); }
};

}  // namespace psd
