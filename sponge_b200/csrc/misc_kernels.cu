// Small sm_100a kernels around the hot path: representation change, synthetic inputs and the K0
// integer-multiply throughput probes.
#include <cuda_runtime.h>

#include "sponge_kernel.cuh"
#include "registry.hpp"

namespace psd {

// ------------------------------------------------------------------------------------------------
// In-place representation change (ark-ff from_bigint / into_bigint), AoS, one element per thread.
// ------------------------------------------------------------------------------------------------
template <class F>
__global__ void __launch_bounds__(256) convert_kernel(PermDesc d, uint32_t* elems, uint64_t n, int to_m,
                                                      int32_t* err) {
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    bool bad = false;
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        uint32_t e[8];
        load_elem(e, elems, kLayoutAos, n, i, 1, 0);
        if (to_m) {
            bad |= !is_canonical<F>(e);
            to_mont<F>(e, d);
        } else {
            from_mont<F>(e);
        }
        store_elem(e, elems, kLayoutAos, n, i, 1, 0);
    }
    if (bad && err) atomicOr(err, 1);
}

// ------------------------------------------------------------------------------------------------
// Synthetic field elements; must match oracle/poseidon_ref.py::synth_element.
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ uint64_t splitmix64(uint64_t x) {
    x += 0x9E3779B97F4A7C15ull;
    uint64_t z = x;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}

template <class F>
__global__ void __launch_bounds__(256) synth_kernel(PermDesc d, uint32_t* elems, uint64_t n, uint64_t seed,
                                                    uint64_t first, int repr) {
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        const uint64_t k = first + i;
        uint32_t e[8];
        for (uint64_t salt = 0;; salt++) {
#pragma unroll
            for (int w = 0; w < 4; w++) {
                const uint64_t ctr = ((seed << 40) ^ (salt << 34)) + 4 * k + w;
                const uint64_t v = splitmix64(ctr);
                e[2 * w] = (uint32_t)v;
                e[2 * w + 1] = (uint32_t)(v >> 32);
            }
            e[7] &= (uint32_t)((1ull << (F::kBits - 224)) - 1);
            if (is_canonical<F>(e)) break;
        }
        if (repr == kReprMontgomery) to_mont<F>(e, d);
        store_elem(e, elems, kLayoutAos, n, i, 1, 0);
    }
}

// ------------------------------------------------------------------------------------------------
// K0: integer-multiply throughput probes.  Every thread runs 8 independent chains and every
// multiply takes one operand from its own running accumulator, so ptxas can neither hoist the
// product out of the loop nor strength-reduce it into adds (a first version with loop-invariant
// operands was silently turned into IADD3 chains).  Results are stored conditionally so nothing is
// optimised away.
//   kind 0: IMAD.WIDE.U32 (32x32+64 -> 64)     kind 1: IMAD (32x32+32 -> low 32)
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) imad_wide_kernel(uint32_t* scratch, uint32_t iters) {
    const uint32_t tid = blockIdx.x * blockDim.x + threadIdx.x;
    const uint32_t b = scratch[(tid + 7) & 1023] | 3u;
    uint64_t acc[8];
#pragma unroll
    for (int k = 0; k < 8; k++) acc[k] = scratch[(tid + k) & 1023] + k;
#pragma unroll 1
    for (uint32_t it = 0; it < iters; it++) {
#pragma unroll
        for (int rep = 0; rep < 8; rep++) {
#pragma unroll
            for (int k = 0; k < 8; k++)
                asm volatile("mad.wide.u32 %0, %1, %2, %0;" : "+l"(acc[k]) : "r"((uint32_t)acc[k]), "r"(b));
        }
    }
    uint64_t s = 0;
#pragma unroll
    for (int k = 0; k < 8; k++) s ^= acc[k];
    if (s == 0x123456789abcdefull) scratch[tid & 1023] = (uint32_t)s;
}

__global__ void __launch_bounds__(256) imad_lo_kernel(uint32_t* scratch, uint32_t iters) {
    const uint32_t tid = blockIdx.x * blockDim.x + threadIdx.x;
    const uint32_t b = scratch[(tid + 7) & 1023] | 3u;
    uint32_t acc[8];
#pragma unroll
    for (int k = 0; k < 8; k++) acc[k] = scratch[(tid + k) & 1023] + k;
#pragma unroll 1
    for (uint32_t it = 0; it < iters; it++) {
#pragma unroll
        for (int rep = 0; rep < 8; rep++) {
#pragma unroll
            for (int k = 0; k < 8; k++) asm volatile("mad.lo.u32 %0, %0, %1, %0;" : "+r"(acc[k]) : "r"(b));
        }
    }
    uint32_t s = 0;
#pragma unroll
    for (int k = 0; k < 8; k++) s ^= acc[k];
    if (s == 0x12345678u) scratch[tid & 1023] = s;
}

// Two independent chains of the library's own Montgomery product (SQR = false) or squaring (true)
// per thread.
template <class F, bool SQR>
__global__ void __launch_bounds__(128) modmul_kernel(uint32_t* scratch, uint32_t iters) {
    const uint32_t tid = blockIdx.x * blockDim.x + threadIdx.x;
    uint32_t x[8], y[8], c[8];
#pragma unroll
    for (int k = 0; k < 8; k++) {
        x[k] = scratch[(tid + k) & 1023];
        y[k] = scratch[(tid + 8 + k) & 1023];
        c[k] = scratch[(16 + k) & 1023];
    }
    x[7] &= 0x0fffffffu;
    y[7] &= 0x0fffffffu;
    c[7] &= 0x0fffffffu;
#pragma unroll 1
    for (uint32_t it = 0; it < iters; it++) {
        if (SQR) {
            fp_sqr<F>(x, x);
            fp_sqr<F>(y, y);
        } else {
            mont_mul<F>(x, x, c);
            mont_mul<F>(y, y, c);
        }
    }
    uint32_t s = 0;
#pragma unroll
    for (int k = 0; k < 8; k++) s ^= x[k] ^ y[k];
    if (s == 0x12345678u) scratch[tid & 1023] = s;
}


// ------------------------------------------------------------------------------------------------
// Launchers
// ------------------------------------------------------------------------------------------------
int launch_convert(int field, const PermDesc& d, uint32_t* elems, uint64_t n, int to_mont_dir, int32_t* err,
                   LaunchCfg cfg, void* stream) {
    if (field == 0)
        convert_kernel<Bn254Fr><<<cfg.grid, cfg.block, 0, (cudaStream_t)stream>>>(d, elems, n, to_mont_dir, err);
    else
        convert_kernel<Bls12381Fr><<<cfg.grid, cfg.block, 0, (cudaStream_t)stream>>>(d, elems, n, to_mont_dir, err);
    return (int)cudaGetLastError();
}

int launch_synth(int field, const PermDesc& d, uint32_t* elems, uint64_t n, uint64_t seed, uint64_t first,
                 int repr, LaunchCfg cfg, void* stream) {
    if (field == 0)
        synth_kernel<Bn254Fr><<<cfg.grid, cfg.block, 0, (cudaStream_t)stream>>>(d, elems, n, seed, first, repr);
    else
        synth_kernel<Bls12381Fr><<<cfg.grid, cfg.block, 0, (cudaStream_t)stream>>>(d, elems, n, seed, first, repr);
    return (int)cudaGetLastError();
}

int launch_imad_peak(int field, int kind, uint32_t* scratch, uint32_t iters, LaunchCfg cfg, void* stream) {
    cudaStream_t s = (cudaStream_t)stream;
    if (kind == 0) imad_wide_kernel<<<cfg.grid, cfg.block, 0, s>>>(scratch, iters);
    else if (kind == 1) imad_lo_kernel<<<cfg.grid, cfg.block, 0, s>>>(scratch, iters);
    else if (kind == 2 && field == 0) modmul_kernel<Bn254Fr, false><<<cfg.grid, cfg.block, 0, s>>>(scratch, iters);
    else if (kind == 2) modmul_kernel<Bls12381Fr, false><<<cfg.grid, cfg.block, 0, s>>>(scratch, iters);
    else if (field == 0) modmul_kernel<Bn254Fr, true><<<cfg.grid, cfg.block, 0, s>>>(scratch, iters);
    else modmul_kernel<Bls12381Fr, true><<<cfg.grid, cfg.block, 0, s>>>(scratch, iters);
    return (int)cudaGetLastError();
}

static int misc_upload(uint32_t word_off, const uint32_t* words, uint32_t n_words, void* stream) {
    return (int)cudaMemcpyToSymbolAsync(c_pool, words, (size_t)n_words * 4, (size_t)word_off * 4,
                                        cudaMemcpyHostToDevice, (cudaStream_t)stream);
}
static ConstPoolRegistrar g_misc_pool(-1, 0, misc_upload);

}  // namespace psd
