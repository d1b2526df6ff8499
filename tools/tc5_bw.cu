// Throughput probes for the variant-7 design (one block of 512 threads = four 128-row tiles, like the kernel):
//   1. tcgen05.ld.32x32b.x32 issued by all 16 warps back to back: TMEM read bytes per cycle per SM;
//   2. tcgen05.mma.kind::i8 M = 128, K = 32 with N = 32 / 64 / 96 / 128 / 256, issued by the four tile leaders to rotating
//      accumulator blocks (no dependent chain): cycles per instruction per SM;
//   3. the same with every MMA accumulating into ONE block (dependent chain).
// Every wait is bounded.  Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/tc5_bw tools/tc5_bw.cu
#include <cuda_runtime.h>

#include <cstdint>
#include <cstdio>

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ uint64_t umma_desc(uint32_t saddr, uint32_t lbo, uint32_t sbo) {
    return (uint64_t)((saddr & 0x3FFFF) >> 4) | ((uint64_t)(lbo >> 4) << 16) | ((uint64_t)(sbo >> 4) << 32) | (1ull << 46);
}
__device__ __forceinline__ uint32_t umma_idesc(uint32_t n) { return (2u << 4) | ((n >> 3) << 17) | ((128u >> 4) << 24); }

__device__ __forceinline__ bool mbar_wait(uint32_t mbar, uint32_t parity) {
    for (uint32_t spin = 0; spin < 20000000u; spin++) {
        uint32_t ok;
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                     : "=r"(ok) : "r"(mbar), "r"(parity) : "memory");
        if (ok) return true;
    }
    return false;
}

__device__ __forceinline__ void ld32(uint32_t taddr, int (&v)[32]) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,"
                 "%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
                 : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]),
                   "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]), "=r"(v[16]),
                   "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), "=r"(v[24]),
                   "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
                 : "r"(taddr) : "memory");
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}

// mode 0: loads; mode 1: MMAs to rotating blocks; mode 2: MMAs into one block.  n = MMA N.  warps = how many warps load.
__global__ void __launch_bounds__(512, 1) probe(int mode, int n, int reps, int warps, long long* cycles, int* err, int* sink) {
    extern __shared__ __align__(1024) unsigned char sm[];
    // A: 4 tiles x 4 KB at 0; B: 8 KB (N up to 256) at 16 KB; mbarriers behind
    __shared__ __align__(8) uint64_t mbar[4];
    __shared__ uint32_t tmem_base;
    const int tid = threadIdx.x, warp = tid >> 5, tile = tid >> 7;
    for (int q = tid; q < (16384 + 8192) / 4; q += 512) reinterpret_cast<uint32_t*>(sm)[q] = 0x01010101u * (q & 3);
    if (tid < 4) asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&mbar[tid])));
    if (tid == 0) asm volatile("fence.mbarrier_init.release.cluster;");
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(smem_u32(&tmem_base)));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    asm volatile("tcgen05.fence::before_thread_sync;");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;");
    const uint32_t tbase = tmem_base;
    int acc = 0;
    long long t0 = clock64();
    if (mode == 0) {
        if (warp < warps) {
            int v[32];
            for (int it = 0; it < reps; it++) {
                const uint32_t taddr = tbase + tile * 128 + (it & 3) * 32 + ((uint32_t)((warp & 3) * 32) << 16);
                ld32(taddr, v);
#pragma unroll
                for (int c = 0; c < 32; c++) acc ^= v[c];
            }
        }
    } else {
        if ((tid & 127) == 0) {
            const uint64_t da = umma_desc(smem_u32(sm) + tile * 4096, 2048, 128);
            const uint64_t db = umma_desc(smem_u32(sm) + 16384, (uint32_t)n * 16, 128);
            const uint32_t idesc = umma_idesc((uint32_t)n);
            const int blocks = mode == 1 ? (128 / n > 0 ? 128 / n : 1) : 1;
            for (int it = 0; it < reps; it++) {
                // tiles own 128 columns each; N = 256 spans two tiles' columns (timing only)
                const uint32_t d = tbase + (n <= 128 ? tile * 128 + (it % blocks) * n : (tile & 1) * 256);
                const uint32_t accf = it >= blocks;
                asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                             "tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, {%5, %5, %5, %5}, p;\n\t}"
                             ::"r"(d), "l"(da), "l"(db), "r"(idesc), "r"(accf), "r"(0u) : "memory");
            }
            asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&mbar[tile])) : "memory");
        }
        if (!mbar_wait(smem_u32(&mbar[tile]), 0)) {
            if ((tid & 127) == 0) *err = 1;
        }
        asm volatile("tcgen05.fence::after_thread_sync;");
    }
    __syncthreads();
    long long t1 = clock64();
    if (acc == 0x7fffffff) *sink = acc;
    if (tid == 0) *cycles = t1 - t0;
    asm volatile("tcgen05.fence::before_thread_sync;");
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tbase));
}

int main() {
    long long* dCyc;
    int *dErr, *dSink;
    cudaMalloc(&dCyc, 8);
    cudaMalloc(&dErr, 4);
    cudaMalloc(&dSink, 4);
    cudaMemset(dErr, 0, 4);
    cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, 32768);
    auto run = [&](int mode, int n, int reps, int warps) {
        probe<<<1, 512, 32768>>>(mode, n, reps, warps, dCyc, dErr, dSink);
        cudaError_t e = cudaDeviceSynchronize();
        long long c = 0;
        int err = 0;
        cudaMemcpy(&c, dCyc, 8, cudaMemcpyDeviceToHost);
        cudaMemcpy(&err, dErr, 4, cudaMemcpyDeviceToHost);
        if (e != cudaSuccess || err) {
            printf("mode %d n %d: cuda=%s wait_timeout=%d\n", mode, n, cudaGetErrorString(e), err);
            exit(1);
        }
        return c;
    };
    for (int warps : {1, 4, 8, 16}) {
        const int reps = 4000;
        run(0, 0, 100, warps);
        const long long c = run(0, 0, reps, warps);
        printf("tcgen05.ld 32x32b.x32 + wait, %2d warps: %.1f cycles per load per warp, %.1f B/cycle/SM\n", warps, (double)c / reps,
               (double)warps * reps * 4096.0 / c);
    }
    for (int mode : {1, 2})
        for (int n : {32, 64, 96, 128, 256}) {
            const int reps = 2000;
            run(mode, n, 50, 0);
            const long long c = run(mode, n, reps, 0);
            printf("tcgen05.mma i8 M128 N%-3d K32, 4 issuing threads, %s: %.1f cycles per MMA per SM\n", n,
                   mode == 1 ? "rotating accumulators" : "one accumulator     ", (double)c / (4.0 * reps));
        }
    printf("status: ok\n");
    return 0;
}
