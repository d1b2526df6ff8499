// Integer/FP64 pipe micro-benchmarks for sm_100a, second pass: operands depend on the running
// accumulators so ptxas cannot hoist or strength-reduce the multiplies.
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <vector>

#define CH 8
#define REPS 16

#define LOOP_BEGIN \
    _Pragma("unroll 1") for (uint32_t it = 0; it < iters; it++) { \
    _Pragma("unroll") for (int r = 0; r < REPS; r++) { \
    _Pragma("unroll") for (int k = 0; k < CH; k++) {
#define LOOP_END }}}

// IMAD.WIDE.U32 acc = lo(acc)*b + acc : dependent chain per accumulator, CH chains per thread
__global__ void k_wide(uint32_t* s, uint32_t iters) {
    uint32_t tid = blockIdx.x * blockDim.x + threadIdx.x;
    uint32_t b = s[(tid + 7) & 1023] | 3u;
    uint64_t acc[CH];
    for (int k = 0; k < CH; k++) acc[k] = s[(tid + k) & 1023] + k;
    LOOP_BEGIN
        asm volatile("mad.wide.u32 %0, %1, %2, %0;" : "+l"(acc[k]) : "r"((uint32_t)acc[k]), "r"(b));
    LOOP_END
    uint64_t x = 0;
    for (int k = 0; k < CH; k++) x ^= acc[k];
    if (x == 0x1234567) s[0] = (uint32_t)x;
}
// IMAD (lo): acc = acc*b + acc
__global__ void k_lo(uint32_t* s, uint32_t iters) {
    uint32_t tid = blockIdx.x * blockDim.x + threadIdx.x;
    uint32_t b = s[(tid + 7) & 1023] | 3u;
    uint32_t acc[CH];
    for (int k = 0; k < CH; k++) acc[k] = s[(tid + k) & 1023] + k;
    LOOP_BEGIN
        asm volatile("mad.lo.u32 %0, %0, %1, %0;" : "+r"(acc[k]) : "r"(b));
    LOOP_END
    uint32_t x = 0;
    for (int k = 0; k < CH; k++) x ^= acc[k];
    if (x == 0x1234567) s[0] = x;
}
// IMAD.HI: acc = hi(acc*b) + acc
__global__ void k_hi(uint32_t* s, uint32_t iters) {
    uint32_t tid = blockIdx.x * blockDim.x + threadIdx.x;
    uint32_t b = s[(tid + 7) & 1023] | 0x80000003u;
    uint32_t acc[CH];
    for (int k = 0; k < CH; k++) acc[k] = s[(tid + k) & 1023] + k;
    LOOP_BEGIN
        asm volatile("mad.hi.u32 %0, %0, %1, %0;" : "+r"(acc[k]) : "r"(b));
    LOOP_END
    uint32_t x = 0;
    for (int k = 0; k < CH; k++) x ^= acc[k];
    if (x == 0x1234567) s[0] = x;
}
// carry chain: (lo,hi) += lo*b with carry-out into top (IMAD.WIDE P-out + IADD3.X)
__global__ void k_wide_cout(uint32_t* s, uint32_t iters) {
    uint32_t tid = blockIdx.x * blockDim.x + threadIdx.x;
    uint32_t b = s[(tid + 7) & 1023] | 3u;
    uint32_t lo[CH], hi[CH], top[CH];
    for (int k = 0; k < CH; k++) { lo[k] = s[(tid + k) & 1023]; hi[k] = tid ^ k; top[k] = 0; }
    LOOP_BEGIN
        asm volatile("mad.lo.cc.u32 %0, %0, %3, %0;\n\tmadc.hi.cc.u32 %1, %0, %3, %1;\n\taddc.u32 %2, %2, 0;"
                     : "+r"(lo[k]), "+r"(hi[k]), "+r"(top[k]) : "r"(b));
    LOOP_END
    uint32_t x = 0;
    for (int k = 0; k < CH; k++) x ^= lo[k] ^ hi[k] ^ top[k];
    if (x == 0x1234567) s[0] = x;
}
// IMAD.WIDE + NADD dependent-but-unmergeable ALU ops (xor/add alternating)
template <int NADD>
__global__ void k_wide_alu(uint32_t* s, uint32_t iters) {
    uint32_t tid = blockIdx.x * blockDim.x + threadIdx.x;
    uint32_t b = s[(tid + 7) & 1023] | 3u;
    uint64_t acc[CH];
    uint32_t z[CH], y[CH];
    for (int k = 0; k < CH; k++) { acc[k] = s[(tid + k) & 1023] + k; z[k] = tid * k; y[k] = tid + k; }
    LOOP_BEGIN
        asm volatile("mad.wide.u32 %0, %1, %2, %0;" : "+l"(acc[k]) : "r"((uint32_t)acc[k]), "r"(b));
#pragma unroll
        for (int q = 0; q < NADD; q++) {
            if (q & 1) asm volatile("add.u32 %0, %0, %1;" : "+r"(z[k]) : "r"(y[k]));
            else asm volatile("xor.b32 %0, %0, %1;" : "+r"(y[k]) : "r"(z[k]));
        }
    LOOP_END
    uint64_t x = 0;
    for (int k = 0; k < CH; k++) x ^= acc[k] ^ z[k] ^ y[k];
    if (x == 0x1234567) s[0] = (uint32_t)x;
}
// ALU only (fibonacci add/xor)
__global__ void k_alu(uint32_t* s, uint32_t iters) {
    uint32_t tid = blockIdx.x * blockDim.x + threadIdx.x;
    uint32_t z[CH], y[CH];
    for (int k = 0; k < CH; k++) { z[k] = s[(tid + k) & 1023]; y[k] = tid + k; }
    LOOP_BEGIN
        asm volatile("add.u32 %0, %0, %1;" : "+r"(z[k]) : "r"(y[k]));
        asm volatile("xor.b32 %0, %0, %1;" : "+r"(y[k]) : "r"(z[k]));
    LOOP_END
    uint32_t x = 0;
    for (int k = 0; k < CH; k++) x ^= z[k] ^ y[k];
    if (x == 0x1234567) s[0] = x;
}
// DFMA dependent chains
__global__ void k_dfma(uint32_t* s, uint32_t iters) {
    uint32_t tid = blockIdx.x * blockDim.x + threadIdx.x;
    double b = 1.0 - s[(tid + 7) & 1023] * 1e-12;
    double acc[CH];
    for (int k = 0; k < CH; k++) acc[k] = 1.0 + (s[(tid + k) & 1023] & 1023) * 1e-6;
    LOOP_BEGIN
        asm volatile("fma.rz.f64 %0, %0, %1, %0;" : "+d"(acc[k]) : "d"(b));
    LOOP_END
    double x = 0;
    for (int k = 0; k < CH; k++) x += acc[k];
    if (x == 0.1234567) s[0] = (uint32_t)x;
}
// DFMA + IMAD.WIDE interleaved 1:1 (separate pipes?)
__global__ void k_dfma_wide(uint32_t* s, uint32_t iters) {
    uint32_t tid = blockIdx.x * blockDim.x + threadIdx.x;
    double bd = 1.0 - s[(tid + 7) & 1023] * 1e-12;
    uint32_t b = s[(tid + 7) & 1023] | 3u;
    double accd[CH];
    uint64_t acc[CH];
    for (int k = 0; k < CH; k++) { accd[k] = 1.0 + (s[(tid + k) & 1023] & 1023) * 1e-6; acc[k] = s[(tid + k) & 1023] + k; }
    LOOP_BEGIN
        asm volatile("fma.rz.f64 %0, %0, %1, %0;" : "+d"(accd[k]) : "d"(bd));
        asm volatile("mad.wide.u32 %0, %1, %2, %0;" : "+l"(acc[k]) : "r"((uint32_t)acc[k]), "r"(b));
    LOOP_END
    double x = 0;
    for (int k = 0; k < CH; k++) x += accd[k] + (double)acc[k];
    if (x == 0.1234567) s[0] = (uint32_t)x;
}
// FFMA + IMAD.WIDE interleaved 1:1 (fmalite vs fmaheavy)
__global__ void k_ffma_wide(uint32_t* s, uint32_t iters) {
    uint32_t tid = blockIdx.x * blockDim.x + threadIdx.x;
    float bf = 1.0f - s[(tid + 7) & 1023] * 1e-12f;
    uint32_t b = s[(tid + 7) & 1023] | 3u;
    float accf[CH];
    uint64_t acc[CH];
    for (int k = 0; k < CH; k++) { accf[k] = 1.0f + (s[(tid + k) & 1023] & 1023) * 1e-6f; acc[k] = s[(tid + k) & 1023] + k; }
    LOOP_BEGIN
        asm volatile("fma.rz.f32 %0, %0, %1, %0;" : "+f"(accf[k]) : "f"(bf));
        asm volatile("mad.wide.u32 %0, %1, %2, %0;" : "+l"(acc[k]) : "r"((uint32_t)acc[k]), "r"(b));
    LOOP_END
    float x = 0;
    for (int k = 0; k < CH; k++) x += accf[k] + (float)acc[k];
    if (x == 0.1234567f) s[0] = (uint32_t)x;
}
// FFMA only
__global__ void k_ffma(uint32_t* s, uint32_t iters) {
    uint32_t tid = blockIdx.x * blockDim.x + threadIdx.x;
    float bf = 1.0f - s[(tid + 7) & 1023] * 1e-12f;
    float accf[CH];
    for (int k = 0; k < CH; k++) accf[k] = 1.0f + (s[(tid + k) & 1023] & 1023) * 1e-6f;
    LOOP_BEGIN
        asm volatile("fma.rz.f32 %0, %0, %1, %0;" : "+f"(accf[k]) : "f"(bf));
    LOOP_END
    float x = 0;
    for (int k = 0; k < CH; k++) x += accf[k];
    if (x == 0.1234567f) s[0] = (uint32_t)x;
}

template <class K>
static void run(const char* name, K kern, double inst_per_thread_iter, int block, int bps, uint32_t iters, uint32_t* d, int sms, double clk) {
    int grid = sms * bps;
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    kern<<<grid, block>>>(d, iters);
    kern<<<grid, block>>>(d, iters);
    cudaEventRecord(e0);
    for (int r = 0; r < 3; r++) kern<<<grid, block>>>(d, iters);
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1); ms /= 3;
    double per_s = (double)grid * block * iters * inst_per_thread_iter / (ms * 1e-3);
    printf("%-26s warps/SMSP=%2d %8.3f ms  %.4e counted-inst/s  %.3f warp-inst/clk/SM  err=%s\n", name, block * bps / 128, ms, per_s,
           per_s / 32.0 / sms / (clk * 1e6), cudaGetErrorString(cudaGetLastError()));
}

int main() {
    cudaDeviceProp prop; cudaGetDeviceProperties(&prop, 0);
    int sms = prop.multiProcessorCount;
    int clk_khz = 0; cudaDeviceGetAttribute(&clk_khz, cudaDevAttrClockRate, 0);
    double clk = clk_khz / 1000.0;
    printf("device %s, %d SMs, clock attr %.0f MHz (rates below assume this clock)\n", prop.name, sms, clk);
    uint32_t* d; cudaMalloc(&d, 4096);
    std::vector<uint32_t> h(1024); for (int i = 0; i < 1024; i++) h[i] = 0x9e3779b9u * (i + 1) + 12345u;
    cudaMemcpy(d, h.data(), 4096, cudaMemcpyHostToDevice);
    const uint32_t it = 2048;
    const double n = CH * REPS;
    for (int bps : {1, 2, 4, 8}) run("IMAD.WIDE", k_wide, n, 256, bps, it, d, sms, clk);
    for (int bps : {2, 8}) run("IMAD lo", k_lo, n, 256, bps, it, d, sms, clk);
    for (int bps : {2, 8}) run("IMAD.HI", k_hi, n, 256, bps, it, d, sms, clk);
    run("IMAD.WIDE c-out +IADD3.X", k_wide_cout, n, 256, 8, it, d, sms, clk);
    run("IMAD.WIDE + 1 ALU", k_wide_alu<1>, n, 256, 8, it, d, sms, clk);
    run("IMAD.WIDE + 2 ALU", k_wide_alu<2>, n, 256, 8, it, d, sms, clk);
    run("IMAD.WIDE + 3 ALU", k_wide_alu<3>, n, 256, 8, it, d, sms, clk);
    run("IMAD.WIDE + 4 ALU", k_wide_alu<4>, n, 256, 8, it, d, sms, clk);
    run("ALU only (2 per)", k_alu, 2 * n, 256, 8, it, d, sms, clk);
    run("DFMA", k_dfma, n, 256, 8, it, d, sms, clk);
    run("DFMA+IMAD.WIDE (pairs)", k_dfma_wide, n, 256, 8, it, d, sms, clk);
    run("FFMA", k_ffma, n, 256, 8, it, d, sms, clk);
    run("FFMA+IMAD.WIDE (pairs)", k_ffma_wide, n, 256, 8, it, d, sms, clk);
    return 0;
}
