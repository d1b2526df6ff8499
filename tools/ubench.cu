// Integer-pipe micro-benchmarks for sm_100a (development tool; results recorded in profiles/).
// Each kernel runs many independent chains per thread; reports warp-instructions/clk/SM.
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <vector>
#include <string>

#define CHAINS 8
#define REPS 8

// 1. plain 64-bit accumulate: IMAD.WIDE.U32 R, a, b, R
__global__ void k_wide_plain(uint32_t* s, uint32_t iters) {
    uint32_t tid = blockIdx.x * blockDim.x + threadIdx.x;
    uint32_t a = s[tid & 1023] | 1u, b = s[(tid + 7) & 1023] | 3u;
    uint64_t acc[CHAINS];
    for (int k = 0; k < CHAINS; k++) acc[k] = tid + k;
#pragma unroll 1
    for (uint32_t it = 0; it < iters; it++)
#pragma unroll
        for (int r = 0; r < REPS; r++)
#pragma unroll
            for (int k = 0; k < CHAINS; k++) asm volatile("mad.wide.u32 %0, %1, %2, %0;" : "+l"(acc[k]) : "r"(a), "r"(b));
    uint64_t x = 0;
    for (int k = 0; k < CHAINS; k++) x ^= acc[k];
    if (x == 0x1234567) s[0] = (uint32_t)x;
}

// 2. carry chain of 4 IMAD.WIDE (first without carry-in, 3 with .X), + addc : the library's mad4
__global__ void k_mad4_chain(uint32_t* s, uint32_t iters) {
    uint32_t tid = blockIdx.x * blockDim.x + threadIdx.x;
    uint32_t a0 = s[tid & 1023] | 1u, a1 = s[(tid + 3) & 1023], a2 = s[(tid + 5) & 1023], a3 = s[(tid + 9) & 1023];
    uint32_t b = s[(tid + 7) & 1023] | 3u;
    uint32_t acc[4][9];
    for (int c = 0; c < 4; c++) for (int k = 0; k < 9; k++) acc[c][k] = tid + k + c;
#pragma unroll 1
    for (uint32_t it = 0; it < iters; it++)
#pragma unroll
        for (int r = 0; r < REPS; r++)
#pragma unroll
            for (int c = 0; c < 4; c++)
                asm volatile("mad.lo.cc.u32 %0, %9, %13, %0;\n\t"
                    "madc.hi.cc.u32 %1, %9, %13, %1;\n\t"
                    "madc.lo.cc.u32 %2, %10, %13, %2;\n\t"
                    "madc.hi.cc.u32 %3, %10, %13, %3;\n\t"
                    "madc.lo.cc.u32 %4, %11, %13, %4;\n\t"
                    "madc.hi.cc.u32 %5, %11, %13, %5;\n\t"
                    "madc.lo.cc.u32 %6, %12, %13, %6;\n\t"
                    "madc.hi.cc.u32 %7, %12, %13, %7;\n\t"
                    "addc.u32 %8, %8, 0;"
                    : "+r"(acc[c][0]), "+r"(acc[c][1]), "+r"(acc[c][2]), "+r"(acc[c][3]), "+r"(acc[c][4]), "+r"(acc[c][5]),
                      "+r"(acc[c][6]), "+r"(acc[c][7]), "+r"(acc[c][8])
                    : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b));
    uint32_t x = 0;
    for (int c = 0; c < 4; c++) for (int k = 0; k < 9; k++) x ^= acc[c][k];
    if (x == 0x1234567) s[0] = x;
}

// 3. carry-OUT only: IMAD.WIDE.U32 R, P, a, b, R ; IADD3.X top += P   (no .X multiplies)
__global__ void k_wide_cout(uint32_t* s, uint32_t iters) {
    uint32_t tid = blockIdx.x * blockDim.x + threadIdx.x;
    uint32_t a = s[tid & 1023] | 1u, b = s[(tid + 7) & 1023] | 3u;
    uint32_t lo[CHAINS], hi[CHAINS], top[CHAINS];
    for (int k = 0; k < CHAINS; k++) { lo[k] = tid + k; hi[k] = tid ^ k; top[k] = 0; }
#pragma unroll 1
    for (uint32_t it = 0; it < iters; it++)
#pragma unroll
        for (int r = 0; r < REPS; r++)
#pragma unroll
            for (int k = 0; k < CHAINS; k++)
                asm volatile("mad.lo.cc.u32 %0, %3, %4, %0;\n\tmadc.hi.cc.u32 %1, %3, %4, %1;\n\taddc.u32 %2, %2, 0;"
                             : "+r"(lo[k]), "+r"(hi[k]), "+r"(top[k]) : "r"(a), "r"(b));
    uint32_t x = 0;
    for (int k = 0; k < CHAINS; k++) x ^= lo[k] ^ hi[k] ^ top[k];
    if (x == 0x1234567) s[0] = x;
}

// 4. plain IMAD.WIDE + N independent IADD3 per multiply (dual-issue test)
template <int NADD>
__global__ void k_wide_plus_add(uint32_t* s, uint32_t iters) {
    uint32_t tid = blockIdx.x * blockDim.x + threadIdx.x;
    uint32_t a = s[tid & 1023] | 1u, b = s[(tid + 7) & 1023] | 3u;
    uint64_t acc[CHAINS];
    uint32_t z[CHAINS];
    for (int k = 0; k < CHAINS; k++) { acc[k] = tid + k; z[k] = tid * k; }
#pragma unroll 1
    for (uint32_t it = 0; it < iters; it++)
#pragma unroll
        for (int r = 0; r < REPS; r++)
#pragma unroll
            for (int k = 0; k < CHAINS; k++) {
                asm volatile("mad.wide.u32 %0, %1, %2, %0;" : "+l"(acc[k]) : "r"(a), "r"(b));
#pragma unroll
                for (int q = 0; q < NADD; q++) asm volatile("add.u32 %0, %0, %1;" : "+r"(z[k]) : "r"(b));
            }
    uint64_t x = 0;
    for (int k = 0; k < CHAINS; k++) x ^= acc[k] ^ z[k];
    if (x == 0x1234567) s[0] = (uint32_t)x;
}

// 5. IADD3 only
__global__ void k_add_only(uint32_t* s, uint32_t iters) {
    uint32_t tid = blockIdx.x * blockDim.x + threadIdx.x;
    uint32_t b = s[(tid + 7) & 1023] | 3u;
    uint32_t z[CHAINS];
    for (int k = 0; k < CHAINS; k++) z[k] = tid * k;
#pragma unroll 1
    for (uint32_t it = 0; it < iters; it++)
#pragma unroll
        for (int r = 0; r < REPS; r++)
#pragma unroll
            for (int k = 0; k < CHAINS; k++) asm volatile("add.u32 %0, %0, %1;" : "+r"(z[k]) : "r"(b));
    uint32_t x = 0;
    for (int k = 0; k < CHAINS; k++) x ^= z[k];
    if (x == 0x1234567) s[0] = x;
}

// 6. add.cc/addc chains (IADD3 with carry predicates)
__global__ void k_addc_chain(uint32_t* s, uint32_t iters) {
    uint32_t tid = blockIdx.x * blockDim.x + threadIdx.x;
    uint32_t b = s[(tid + 7) & 1023] | 3u;
    uint32_t z[CHAINS][2];
    for (int k = 0; k < CHAINS; k++) { z[k][0] = tid * k; z[k][1] = tid + k; }
#pragma unroll 1
    for (uint32_t it = 0; it < iters; it++)
#pragma unroll
        for (int r = 0; r < REPS; r++)
#pragma unroll
            for (int k = 0; k < CHAINS; k++)
                asm volatile("add.cc.u32 %0, %0, %2;\n\taddc.u32 %1, %1, %2;" : "+r"(z[k][0]), "+r"(z[k][1]) : "r"(b));
    uint32_t x = 0;
    for (int k = 0; k < CHAINS; k++) x ^= z[k][0] ^ z[k][1];
    if (x == 0x1234567) s[0] = x;
}

// 7. 32-bit IMAD lo with 64-bit shift (funnel) mix: IMAD + SHF
__global__ void k_wide_plus_shf(uint32_t* s, uint32_t iters) {
    uint32_t tid = blockIdx.x * blockDim.x + threadIdx.x;
    uint32_t a = s[tid & 1023] | 1u, b = s[(tid + 7) & 1023] | 3u;
    uint64_t acc[CHAINS];
    uint32_t z[CHAINS];
    for (int k = 0; k < CHAINS; k++) { acc[k] = tid + k; z[k] = tid * k; }
#pragma unroll 1
    for (uint32_t it = 0; it < iters; it++)
#pragma unroll
        for (int r = 0; r < REPS; r++)
#pragma unroll
            for (int k = 0; k < CHAINS; k++) {
                asm volatile("mad.wide.u32 %0, %1, %2, %0;" : "+l"(acc[k]) : "r"(a), "r"(b));
                asm volatile("shf.r.wrap.b32 %0, %0, %1, 29;" : "+r"(z[k]) : "r"(b));
            }
    uint64_t x = 0;
    for (int k = 0; k < CHAINS; k++) x ^= acc[k] ^ z[k];
    if (x == 0x1234567) s[0] = (uint32_t)x;
}

// 8. IMAD.WIDE with uniform-register / constant operand b (c[bank] operand)
__constant__ uint32_t cK[64];
__global__ void k_wide_const(uint32_t* s, uint32_t iters) {
    uint32_t tid = blockIdx.x * blockDim.x + threadIdx.x;
    uint32_t a = s[tid & 1023] | 1u;
    uint64_t acc[CHAINS];
    for (int k = 0; k < CHAINS; k++) acc[k] = tid + k;
#pragma unroll 1
    for (uint32_t it = 0; it < iters; it++)
#pragma unroll
        for (int r = 0; r < REPS; r++)
#pragma unroll
            for (int k = 0; k < CHAINS; k++) asm volatile("mad.wide.u32 %0, %1, %2, %0;" : "+l"(acc[k]) : "r"(a), "r"(cK[(it & 7) * 8 + k]));
    uint64_t x = 0;
    for (int k = 0; k < CHAINS; k++) x ^= acc[k];
    if (x == 0x1234567) s[0] = (uint32_t)x;
}

// 9. FP64 DFMA rate (for reference)
__global__ void k_dfma(uint32_t* s, uint32_t iters) {
    uint32_t tid = blockIdx.x * blockDim.x + threadIdx.x;
    double a = 1.0 + s[tid & 1023] * 1e-9, b = 1.0 - s[(tid + 7) & 1023] * 1e-9;
    double acc[CHAINS];
    for (int k = 0; k < CHAINS; k++) acc[k] = tid + k;
#pragma unroll 1
    for (uint32_t it = 0; it < iters; it++)
#pragma unroll
        for (int r = 0; r < REPS; r++)
#pragma unroll
            for (int k = 0; k < CHAINS; k++) asm volatile("fma.rn.f64 %0, %1, %2, %0;" : "+d"(acc[k]) : "d"(a), "d"(b));
    double x = 0;
    for (int k = 0; k < CHAINS; k++) x += acc[k];
    if (x == 0.1234567) s[0] = (uint32_t)x;
}

template <class K>
static void run(const char* name, K kern, double inst_per_thread_iter, int block, int blocks_per_sm, uint32_t iters, uint32_t* d, int sms, double clk_mhz) {
    int grid = sms * blocks_per_sm;
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    kern<<<grid, block>>>(d, iters);
    kern<<<grid, block>>>(d, iters);
    cudaEventRecord(e0);
    for (int r = 0; r < 3; r++) kern<<<grid, block>>>(d, iters);
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1); ms /= 3;
    double thread_inst = (double)grid * block * iters * inst_per_thread_iter;
    double per_s = thread_inst / (ms * 1e-3);
    double warp_inst_per_clk_sm = per_s / 32.0 / sms / (clk_mhz * 1e6);
    printf("%-28s block=%d x%d  %8.3f ms  %.4e thread-inst/s  %.3f warp-inst/clk/SM (at %.0f MHz)  err=%s\n", name, block, blocks_per_sm, ms, per_s,
           warp_inst_per_clk_sm, clk_mhz, cudaGetErrorString(cudaGetLastError()));
}

int main() {
    cudaDeviceProp prop; cudaGetDeviceProperties(&prop, 0);
    int sms = prop.multiProcessorCount;
    int clk_khz = 0; cudaDeviceGetAttribute(&clk_khz, cudaDevAttrClockRate, 0);
    double clk = clk_khz / 1000.0;
    printf("device %s, %d SMs, clock attr %.0f MHz\n", prop.name, sms, clk);
    uint32_t* d; cudaMalloc(&d, 4096);
    std::vector<uint32_t> h(1024); for (int i = 0; i < 1024; i++) h[i] = 0x9e3779b9u * (i + 1) + 12345u;
    cudaMemcpy(d, h.data(), 4096, cudaMemcpyHostToDevice);
    cudaMemcpyToSymbol(cK, h.data(), 256);
    const uint32_t it = 4096;
    for (int bps : {2, 4, 8}) {
        run("wide_plain", k_wide_plain, CHAINS * REPS, 256, bps, it, d, sms, clk);
    }
    run("mad4_chain (IMAD count)", k_mad4_chain, 4 * 4 * REPS, 256, 4, it, d, sms, clk);
    run("mad4_chain (IMAD count)", k_mad4_chain, 4 * 4 * REPS, 256, 8, it, d, sms, clk);
    run("wide_cout (IMAD count)", k_wide_cout, CHAINS * REPS, 256, 8, it, d, sms, clk);
    run("wide+1add (IMAD count)", k_wide_plus_add<1>, CHAINS * REPS, 256, 8, it, d, sms, clk);
    run("wide+2add (IMAD count)", k_wide_plus_add<2>, CHAINS * REPS, 256, 8, it, d, sms, clk);
    run("wide+3add (IMAD count)", k_wide_plus_add<3>, CHAINS * REPS, 256, 8, it, d, sms, clk);
    run("add_only", k_add_only, CHAINS * REPS, 256, 8, it, d, sms, clk);
    run("addc_chain (2 per)", k_addc_chain, 2 * CHAINS * REPS, 256, 8, it, d, sms, clk);
    run("wide+shf (IMAD count)", k_wide_plus_shf, CHAINS * REPS, 256, 8, it, d, sms, clk);
    run("wide_const", k_wide_const, CHAINS * REPS, 256, 8, it, d, sms, clk);
    run("dfma", k_dfma, CHAINS * REPS, 256, 8, it, d, sms, clk);
    return 0;
}
