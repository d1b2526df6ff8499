// Micro-benchmarks for the variant-5 building blocks on sm_100a: what one SM sub-partition pays, in
// cycles per warp, for (a) 25 table products on the FP64 pipe incl. their integer accumulation,
// (b) the same FP64 instructions alone, (c) the integer accumulation alone, (d) one integer S-box,
// (e) S-box and table products in one basic block.  Build: nvcc -arch=sm_100a -O3 -o tools/ubench3 tools/ubench3.cu
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <vector>

#include "../sponge_b200/csrc/poseidon_fp52.cuh"

using namespace psd;

#define TAB_WORDS 2000   // doubles of pool used as rolling table area

__device__ __forceinline__ double opaque104() {
    return __hiloint2double((int)(0x46700000u + (__activemask() == 0u ? 1u : 0u)), 0);
}

// (a) 25 products + accumulation per iteration.  MODE 0: table at a fixed address (constant-bank
// immediates), 1: address advances with the loop counter (uniform loads), 2: wraps with % (what ptxas
// turned into per-thread LDC in the first version of this test)
template <int MODE>
__global__ void __launch_bounds__(128) k_fp_full(uint64_t* out, uint32_t iters) {
    const double c104 = opaque104();
    double a[5];
    for (int i = 0; i < 5; i++) a[i] = (double)((threadIdx.x * 977u + i * 131u) & 0xfffff);
    uint64_t c[6] = {0, 0, 0, 0, 0, 0};
    uint32_t off = 0;
#pragma unroll 1
    for (uint32_t it = 0; it < iters; it++) {
        const double* tab = f52_pool_d(MODE == 0 ? 0u : MODE == 1 ? (it & 63u) * 25u : off);
#pragma unroll
        for (int i = 0; i < 5; i++) f52_mac5(c, a[i], tab + i * 5, c104);
        off = (off + 25) % TAB_WORDS;
    }
    uint64_t x = 0;
    for (int k = 0; k < 6; k++) x ^= c[k];
    if (x == 0x123456789ull) out[0] = x;
}

// (b) uniform 64-bit constant loads only: 25 per iteration, folded with xor (13 LOP3 pairs)
__global__ void __launch_bounds__(128) k_ldcu(uint64_t* out, uint32_t iters) {
    uint64_t x = threadIdx.x;
#pragma unroll 1
    for (uint32_t it = 0; it < iters; it++) {
        const uint64_t* tab = f52_pool_u((it & 63u) * 25u);
        uint64_t y = 0;
#pragma unroll
        for (int k = 0; k < 25; k++) y ^= tab[k];
        x += y;
    }
    if (x == 0x123456789ull) out[0] = x;
}

// (a2) 25 products with every operand in registers (no constant loads).  ACC 1: integer accumulation
// as in f52_mac5; ACC 0: results folded with FP64 adds instead (100 FP64 instructions, no ALU work).
template <int ACC>
__global__ void __launch_bounds__(128) k_fp_reg(uint64_t* out, uint32_t iters) {
    const double c104 = opaque104();
    double a[5], b[5];
    for (int i = 0; i < 5; i++) {
        a[i] = (double)((threadIdx.x * 977u + i * 131u) & 0xfffff);
        b[i] = (double)(out[(threadIdx.x + i) & 63] >> 13);
    }
    uint64_t c[6] = {0, 0, 0, 0, 0, 0};
    double facc = 0.0;
#pragma unroll 1
    for (uint32_t it = 0; it < iters; it++) {
#pragma unroll
        for (int i = 0; i < 5; i++) {
            uint64_t hi[5], lo[5];
#pragma unroll
            for (int k = 0; k < 5; k++) f52_prod(hi[k], lo[k], a[i], b[k], c104);
            if (ACC) {
                c[0] += lo[0];
#pragma unroll
                for (int k = 1; k < 5; k++) c[k] += lo[k] + hi[k - 1];
                c[5] += hi[4];
            } else {
#pragma unroll
                for (int k = 0; k < 5; k++) facc = __dadd_rz(facc, __longlong_as_double(lo[k]));
            }
        }
        // keep the multiplicands loop-carried: one LOP3 pair each
#pragma unroll
        for (int i = 0; i < 5; i++) a[i] = __longlong_as_double(__double_as_longlong(a[i]) ^ 0x100ll);
    }
    uint64_t x = (uint64_t)__double_as_longlong(facc);
    for (int k = 0; k < 6; k++) x ^= c[k];
    if (x == 0x123456789ull) out[0] = x;
}

// (a3) FP64 instruction mixes, register operands, 100 instructions per iteration in 25 independent
// chains of 4:  MIX 0: fma.rz only; 1: fma.rn only; 2: add.rz only; 3: fma.rz, add.rz alternating;
// 4: fma.rz with the subtraction done as a second fma (hi * -1 + C2)
template <int MIX>
__global__ void __launch_bounds__(128) k_fp64_mix(uint64_t* out, uint32_t iters) {
    double v[25], b = (double)(out[threadIdx.x & 63] >> 13), c2 = 20282409603651674927546878656512.0;
    for (int i = 0; i < 25; i++) v[i] = (double)((threadIdx.x * 977u + i * 131u) & 0xfffff);
#pragma unroll 1
    for (uint32_t it = 0; it < iters; it++) {
#pragma unroll
        for (int i = 0; i < 25; i++) {
            double x = v[i];
#pragma unroll
            for (int q = 0; q < 4; q++) {
                if (MIX == 0) x = __fma_rz(x, b, c2);
                else if (MIX == 1) x = __fma_rn(x, b, c2);
                else if (MIX == 2) x = __dadd_rz(x, b);
                else if (MIX == 3) x = (q & 1) ? __dadd_rz(c2, -x) : __fma_rz(x, b, c2);
                else x = (q & 1) ? __fma_rz(x, -1.0, c2) : __fma_rz(x, b, c2);
            }
            v[i] = x;
        }
    }
    double s = 0;
    for (int i = 0; i < 25; i++) s += v[i];
    if (s == 1.2345) out[0] = 1;
}

// (c) the accumulation of (a) only: 25 three-input 64-bit adds per iteration on register values
__global__ void __launch_bounds__(128) k_alu_only(uint64_t* out, uint32_t iters) {
    uint64_t v[10];
    for (int i = 0; i < 10; i++) v[i] = out[(threadIdx.x + i) & 63] | 1;
    uint64_t c[6] = {0, 0, 0, 0, 0, 0};
#pragma unroll 1
    for (uint32_t it = 0; it < iters; it++) {
#pragma unroll
        for (int i = 0; i < 5; i++) {
            c[0] += v[i];
#pragma unroll
            for (int k = 1; k < 5; k++) c[k] += v[(i + k) % 10] + v[(i + k + 5) % 10];
            c[5] += v[i + 5];
        }
#pragma unroll
        for (int i = 0; i < 10; i++) v[i] ^= c[i % 6] >> 7;   // keeps v loop-carried (20 more ALU ops)
    }
    uint64_t x = 0;
    for (int k = 0; k < 6; k++) x ^= c[k];
    if (x == 0x123456789ull) out[0] = x;
}

// (d) one lazy integer S-box (x^5) per iteration
__global__ void __launch_bounds__(128) k_sbox(uint64_t* out, uint32_t iters) {
    uint32_t x[8];
    for (int i = 0; i < 8; i++) x[i] = (uint32_t)out[(threadIdx.x + i) & 63] >> (i == 7 ? 3 : 0);
#pragma unroll 1
    for (uint32_t it = 0; it < iters; it++) {
        f52_sbox_straight<Bn254Fr, 5>(x);
        x[7] &= 0x1fffffffu;
    }
    uint32_t s = 0;
    for (int i = 0; i < 8; i++) s ^= x[i];
    if (s == 0x1234567u) out[0] = s;
}

// (e) S-box and N x 25 independent table products in one basic block
template <int NFP>
__global__ void __launch_bounds__(128) k_sbox_fp(uint64_t* out, uint32_t iters) {
    const double c104 = opaque104();
    uint32_t x[8];
    for (int i = 0; i < 8; i++) x[i] = (uint32_t)out[(threadIdx.x + i) & 63] >> (i == 7 ? 3 : 0);
    double a[5];
    for (int i = 0; i < 5; i++) a[i] = (double)((threadIdx.x * 977u + i * 131u) & 0xfffff);
    uint64_t c[6] = {0, 0, 0, 0, 0, 0};
    uint32_t off = 0;
#pragma unroll 1
    for (uint32_t it = 0; it < iters; it++) {
        f52_sbox_straight<Bn254Fr, 5>(x);
        x[7] &= 0x1fffffffu;
#pragma unroll
        for (int n = 0; n < NFP; n++) {
            const double* tab = f52_pool_d(off + n * 25);
#pragma unroll
            for (int i = 0; i < 5; i++) f52_mac5(c, a[i], tab + i * 5, c104);
        }
        off = (off + 25 * NFP) % TAB_WORDS;
    }
    uint64_t s = 0;
    for (int i = 0; i < 8; i++) s ^= x[i];
    for (int k = 0; k < 6; k++) s ^= c[k];
    if (s == 0x123456789ull) out[0] = s;
}

// (f) cross-warp overlap: blocks of 256 threads (2 warps per SMSP); warps 0-3 run S-boxes, warps 4-7
// run table products (register operands), SPLIT = 0: every warp does both, one after the other
template <int SPLIT>
__global__ void __launch_bounds__(256) k_mixed(uint64_t* out, uint32_t iters) {
    const double c104 = opaque104();
    uint32_t x[8];
    for (int i = 0; i < 8; i++) x[i] = (uint32_t)out[(threadIdx.x + i) & 63] >> (i == 7 ? 3 : 0);
    double a[5], b[5];
    for (int i = 0; i < 5; i++) {
        a[i] = (double)((threadIdx.x * 977u + i * 131u) & 0xfffff);
        b[i] = (double)(out[(threadIdx.x + i) & 63] >> 13);
    }
    uint64_t c[6] = {0, 0, 0, 0, 0, 0};
    const bool do_int = SPLIT ? (threadIdx.x < 128) : true, do_fp = SPLIT ? (threadIdx.x >= 128) : true;
    // per iteration: 2 S-boxes or 8 x 25 products when split, 1 S-box + 4 x 25 products when not:
    // the same total work per block either way
    const int n_int = SPLIT ? 2 : 1, n_fp = SPLIT ? 8 : 4;
#pragma unroll 1
    for (uint32_t it = 0; it < iters; it++) {
        if (do_int) {
#pragma unroll 1
            for (int r = 0; r < n_int; r++) {
                f52_sbox_straight<Bn254Fr, 5>(x);
                x[7] &= 0x1fffffffu;
            }
        }
        if (do_fp) {
#pragma unroll 1
            for (int r = 0; r < n_fp; r++) {
#pragma unroll
                for (int i = 0; i < 5; i++) {
                    uint64_t hi[5], lo[5];
#pragma unroll
                    for (int k = 0; k < 5; k++) f52_prod(hi[k], lo[k], a[i], b[k], c104);
                    c[0] += lo[0];
#pragma unroll
                    for (int k = 1; k < 5; k++) c[k] += lo[k] + hi[k - 1];
                    c[5] += hi[4];
                }
#pragma unroll
                for (int i = 0; i < 5; i++) a[i] = __longlong_as_double(__double_as_longlong(a[i]) ^ 0x100ll);
            }
        }
    }
    uint64_t s = 0;
    for (int i = 0; i < 8; i++) s ^= x[i];
    for (int k = 0; k < 6; k++) s ^= c[k];
    if (s == 0x123456789ull) out[0] = s;
}

template <class K>
static void run256(const char* name, K kern, int blocks_per_sm, uint32_t iters, uint64_t* d, int sms, double clk_mhz) {
    kern<<<sms * blocks_per_sm, 256>>>(d, 16);
    cudaDeviceSynchronize();
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    cudaEventRecord(e0);
    kern<<<sms * blocks_per_sm, 256>>>(d, iters);
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms = 0;
    cudaEventElapsedTime(&ms, e0, e1);
    const double cyc = ms * 1e-3 * clk_mhz * 1e6;
    printf("%-34s blocks/SM=%d  %8.3f ms  SMSP cycles per (1 S-box + 100 products) = %8.1f   [%s]\n", name,
           blocks_per_sm, ms, cyc / iters / blocks_per_sm / 2, cudaGetErrorString(cudaGetLastError()));
}

template <class K>
static void run(const char* name, K kern, int blocks_per_sm, uint32_t iters, uint64_t* d, int sms, double clk_mhz) {
    const int block = 128;   // 1 warp per SMSP per block
    kern<<<sms * blocks_per_sm, block>>>(d, 16);
    cudaDeviceSynchronize();
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    cudaEventRecord(e0);
    kern<<<sms * blocks_per_sm, block>>>(d, iters);
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms = 0;
    cudaEventElapsedTime(&ms, e0, e1);
    const double cyc = ms * 1e-3 * clk_mhz * 1e6;
    printf("%-34s warps/SMSP=%d  %8.3f ms  cycles per iteration per warp (SMSP time / warps) = %8.1f   [%s]\n", name,
           blocks_per_sm, ms, cyc / iters / blocks_per_sm, cudaGetErrorString(cudaGetLastError()));
}

int main() {
    cudaDeviceProp prop;
    cudaGetDeviceProperties(&prop, 0);
    int sms = prop.multiProcessorCount, clk_khz = 0;
    cudaDeviceGetAttribute(&clk_khz, cudaDevAttrClockRate, 0);
    const double clk = clk_khz / 1000.0;
    printf("device %s, %d SMs, clock attr %.0f MHz\n", prop.name, sms, clk);
    std::vector<double> tab(TAB_WORDS + 64);
    for (size_t i = 0; i < tab.size(); i++) tab[i] = (double)((0x9e3779b97f4a7c15ull * (i + 1)) >> 12);
    cudaMemcpyToSymbol(c_pool, tab.data(), tab.size() * 8);
    uint64_t* d;
    cudaMalloc(&d, 4096);
    std::vector<uint64_t> h(512);
    for (int i = 0; i < 512; i++) h[i] = 0x9e3779b97f4a7c15ull * (i + 3);
    cudaMemcpy(d, h.data(), 4096, cudaMemcpyHostToDevice);
    const uint32_t it = 4096;
    for (int w : {4, 8}) run("100 x fma.rz", k_fp64_mix<0>, w, it, d, sms, clk);
    for (int w : {4, 8}) run("100 x fma.rn", k_fp64_mix<1>, w, it, d, sms, clk);
    for (int w : {4, 8}) run("100 x add.rz", k_fp64_mix<2>, w, it, d, sms, clk);
    for (int w : {4, 8}) run("50 fma.rz + 50 add.rz", k_fp64_mix<3>, w, it, d, sms, clk);
    for (int w : {4, 8}) run("100 x fma.rz (sub as fma)", k_fp64_mix<4>, w, it, d, sms, clk);
    for (int w : {1, 2, 4, 6}) run("25 products + int acc, registers", k_fp_reg<1>, w, it, d, sms, clk);
    for (int w : {1, 2, 4, 6}) run("25 products + FP acc, registers", k_fp_reg<0>, w, it, d, sms, clk);
    for (int w : {1, 4, 6}) run("25 products + acc, uniform loads", k_fp_full<1>, w, it, d, sms, clk);
    for (int w : {4}) run("25 products + acc, % offset", k_fp_full<2>, w, it, d, sms, clk);
    for (int w : {1, 4}) run("25 uniform 64-bit loads", k_ldcu, w, it, d, sms, clk);
    for (int w : {1, 4}) run("25x 3-input 64-bit adds (+20 ALU)", k_alu_only, w, it, d, sms, clk);
    for (int w : {1, 2, 4, 6}) run("integer S-box x^5 (lazy)", k_sbox, w, it, d, sms, clk);
    for (int w : {4}) run("S-box + 50 products, one block", k_sbox_fp<2>, w, it, d, sms, clk);
    for (int w : {1, 2, 3}) run256("every warp: S-box then products", k_mixed<0>, w, it, d, sms, clk);
    for (int w : {1, 2, 3}) run256("half the warps each", k_mixed<1>, w, it, d, sms, clk);
    return 0;
}
