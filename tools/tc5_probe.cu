// Probe for the tcgen05 form of variant 6's contraction: ONE 128-sponge tile, A (128 values x 32 bytes) and a byte-digit
// table B (32 columns x 32 bytes) in shared memory in the canonical K-major no-swizzle layout, D = A x B^T accumulated in
// TMEM by tcgen05.mma.kind::i8 (u8 x u8 -> s32, M = 128, N = 32, K = 32 per instruction), read back with
// tcgen05.ld.32x32b so that thread r receives the 32 column sums of row r.  Checks the result against the host and times the
// round trip (write A -> fence -> barrier -> mma -> commit -> mbarrier wait -> ld) that a partial round would pay.
// Every wait is bounded: a wrong descriptor must end in an error code, not in a hung GPU.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/tc5_probe tools/tc5_probe.cu
#include <cuda_runtime.h>

#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <vector>

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

// K-major, no swizzle: 8-row x 16-byte core matrices; lbo = bytes between the two 16-byte K chunks, sbo = bytes between
// 8-row groups
__device__ __forceinline__ uint64_t umma_desc(uint32_t saddr, uint32_t lbo, uint32_t sbo) {
    return (uint64_t)((saddr & 0x3FFFF) >> 4) | ((uint64_t)(lbo >> 4) << 16) | ((uint64_t)(sbo >> 4) << 32) | (1ull << 46);
}

// kind::i8, u8 x u8 -> s32, K-major A and B, M = 128, N = n
__host__ __device__ constexpr uint32_t umma_idesc(uint32_t n) { return (2u << 4) | ((n >> 3) << 17) | ((128u >> 4) << 24); }

__device__ __forceinline__ bool mbar_wait(uint32_t mbar, uint32_t parity) {
    for (uint32_t spin = 0; spin < 20000000u; spin++) {
        uint32_t ok;
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                     : "=r"(ok) : "r"(mbar), "r"(parity) : "memory");
        if (ok) return true;
    }
    return false;
}

__global__ void __launch_bounds__(128, 1) probe(const uint8_t* A_g, const uint8_t* B_g, int* D_g, int ksteps, int reps,
                                                long long* cycles, int* err) {
    __shared__ __align__(128) uint8_t sA[4096];   // [kk][row][16]
    __shared__ __align__(128) uint8_t sB[1024];   // [kk][n][16]
    __shared__ __align__(8) uint64_t mbar;
    __shared__ uint32_t tmem_base;
    const int r = threadIdx.x, warp = r >> 5;
    if (r == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&mbar)));
        asm volatile("fence.mbarrier_init.release.cluster;");
    }
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 32;" ::"r"(smem_u32(&tmem_base)));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
    }
    // operands
    {
        const uint4* src = reinterpret_cast<const uint4*>(A_g + r * 32);
        *reinterpret_cast<uint4*>(sA + r * 16) = src[0];
        *reinterpret_cast<uint4*>(sA + 2048 + r * 16) = src[1];
        if (r < 32) {
            const uint4* b = reinterpret_cast<const uint4*>(B_g + r * 32);
            *reinterpret_cast<uint4*>(sB + r * 16) = b[0];
            *reinterpret_cast<uint4*>(sB + 512 + r * 16) = b[1];
        }
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    asm volatile("tcgen05.fence::before_thread_sync;");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;");
    const uint32_t tbase = tmem_base;
    const uint64_t da = umma_desc(smem_u32(sA), 2048, 128), db = umma_desc(smem_u32(sB), 512, 128);
    const uint32_t idesc = umma_idesc(32);
    int v[32];
    uint32_t parity = 0;
    long long t0 = clock64();
    for (int it = 0; it < reps; it++) {
        if (r == 0) {
            for (int k = 0; k < ksteps; k++) {
                const uint32_t acc = k > 0;
                asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                             "tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, {%5, %5, %5, %5}, p;\n\t}"
                             ::"r"(tbase), "l"(da), "l"(db), "r"(idesc), "r"(acc), "r"(0u) : "memory");
            }
            asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&mbar)) : "memory");
        }
        if (!mbar_wait(smem_u32(&mbar), parity)) {
            if (r == 0) *err = 1;
            break;
        }
        parity ^= 1;
        asm volatile("tcgen05.fence::after_thread_sync;");
        const uint32_t taddr = tbase + ((uint32_t)((warp & 3) * 32) << 16);
        asm volatile("tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,"
                     "%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
                     : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]),
                       "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]), "=r"(v[16]),
                       "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), "=r"(v[24]),
                       "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
                     : "r"(taddr));
        asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
        // the next iteration's first MMA overwrites the accumulator: order the loads before it
        asm volatile("tcgen05.fence::before_thread_sync;");
        __syncthreads();
        asm volatile("tcgen05.fence::after_thread_sync;");
    }
    long long t1 = clock64();
    for (int c = 0; c < 32; c++) D_g[r * 32 + c] = v[c];
    if (r == 0) *cycles = t1 - t0;
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 32;" ::"r"(tbase));
}

int main() {
    std::vector<uint8_t> A(128 * 32), B(32 * 32);
    for (size_t i = 0; i < A.size(); i++) A[i] = (uint8_t)(i * 131u + 7u * (i >> 5));
    for (size_t i = 0; i < B.size(); i++) B[i] = (uint8_t)(i * 197u + 13u);
    uint8_t *dA, *dB;
    int *dD, *dErr;
    long long* dCyc;
    cudaMalloc(&dA, A.size());
    cudaMalloc(&dB, B.size());
    cudaMalloc(&dD, 128 * 32 * 4);
    cudaMalloc(&dErr, 4);
    cudaMalloc(&dCyc, 8);
    cudaMemcpy(dA, A.data(), A.size(), cudaMemcpyHostToDevice);
    cudaMemcpy(dB, B.data(), B.size(), cudaMemcpyHostToDevice);
    int rc = 0;
    for (int ksteps = 1; ksteps <= 2; ksteps++) {
        cudaMemset(dErr, 0, 4);
        cudaMemset(dD, 0xff, 128 * 32 * 4);
        probe<<<1, 128>>>(dA, dB, dD, ksteps, 1, dCyc, dErr);
        cudaError_t e = cudaDeviceSynchronize();
        int err = 0;
        cudaMemcpy(&err, dErr, 4, cudaMemcpyDeviceToHost);
        std::vector<int> D(128 * 32);
        cudaMemcpy(D.data(), dD, D.size() * 4, cudaMemcpyDeviceToHost);
        long bad = 0;
        for (int m = 0; m < 128; m++)
            for (int n = 0; n < 32; n++) {
                int s = 0;
                for (int k = 0; k < 32; k++) s += (int)A[m * 32 + k] * (int)B[n * 32 + k];   // D[m][n] = sum_k A[m][k] B[n][k]
                s *= ksteps;
                if (D[m * 32 + n] != s) {
                    if (bad < 5) printf("  mismatch m=%d n=%d got %d want %d\n", m, n, D[m * 32 + n], s);
                    bad++;
                }
            }
        printf("ksteps=%d: cuda=%s wait_timeout=%d mismatches=%ld of 4096\n", ksteps, cudaGetErrorString(e), err, bad);
        if (e != cudaSuccess || err || bad) rc = 1;
        if (e != cudaSuccess) return 2;
    }
    if (rc == 0) {
        for (int ksteps : {1, 5, 9}) {
            const int reps = 2000;
            probe<<<1, 128>>>(dA, dB, dD, ksteps, reps, dCyc, dErr);
            cudaDeviceSynchronize();
            long long cyc = 0;
            cudaMemcpy(&cyc, dCyc, 8, cudaMemcpyDeviceToHost);
            printf("round trip with %d MMA(s) of 128x32x32 + commit + wait + tcgen05.ld + barrier: %.0f cycles\n", ksteps,
                   (double)cyc / reps);
        }
    }
    printf("status: %s\n", rc ? "FAILED" : "ok");
    return rc;
}
