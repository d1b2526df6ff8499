// Micro-benchmark for VERDICT r1 item 10: can the CONSTANT linear layer of Poseidon run on the tensor cores?
//
// Digit-table form of a product by a round constant c:  x * c * 2^32  ==  sum_k byte_k(x) * Tab_c[k]  (mod p),
// Tab_c[k] = c * 256^k * 2^32 mod p written as 32 bytes.  Over a batch this IS a contraction:
//     Cols[sponge, col] = sum_{lane j, byte k} byte_{j,k}(sponge) * Tab_{c_j}[k][col]       (u8 x u8 -> s32)
// an (M = sponges) x (K = 32 per lane) x (N = 32 columns per output value) integer GEMM whose column sums stay
// below 2^24, followed per output value by a carry sweep (32 byte-weighted columns -> 9 x 32-bit limbs) and ONE
// 32-bit Montgomery row.  Here: mma.sync.m16n8k32 (u8 x u8 -> s32, register fragments), A fragments built from
// the thread-per-sponge registers through shared memory + ldmatrix, B fragments (tables) from shared memory,
// column sums brought back to the owning thread through shared memory.
//
// Kernels (512-thread blocks, one per SM, as the shipped variant-5 kernel):
//   k_imma      raw IMMA issue rate
//   k_round<0>  the integer lazy S-box x^5 alone (the part that stays on the integer multiplier)
//   k_round<1>  one mock PARTIAL round, t = 3:  S-box; row 0 over (y, lane1, lane2); two lane updates; three
//               sweeps; lanes re-packed to A fragments -- the same data flow and dependency chain as the real round
//   k_round<2>  as <1> without the S-box (the linear layer alone)
// Output: cycles per warp-round per SM sub-partition, to set against the shipped kernel's 2 833 (variant 5) and
// 3 290 (variant 3) cycles per partial round.  Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/ubench4 tools/ubench4.cu
#include <cuda_runtime.h>

#include <cstdint>
#include <cstdio>
#include <vector>

#include "../sponge_b200/csrc/poseidon_core.cuh"

using namespace psd;
typedef Bn254Fr F;

__device__ __forceinline__ void mma_u8(int (&d)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
    asm volatile("mma.sync.aligned.m16n8k32.row.col.s32.u8.u8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+r"(d[0]), "+r"(d[1]), "+r"(d[2]), "+r"(d[3])
                 : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}

__device__ __forceinline__ void ldmatrix_x4(uint32_t (&r)[4], uint32_t smem_addr) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(smem_addr));
}

__global__ void __launch_bounds__(512, 1) k_imma(int* out, uint32_t iters) {
    int acc[8][4];
    uint32_t a[4], b0 = threadIdx.x * 2654435761u, b1 = b0 ^ 0x9e3779b9u;
#pragma unroll
    for (int i = 0; i < 4; i++) a[i] = threadIdx.x * 40503u + i * 977u;
#pragma unroll
    for (int q = 0; q < 8; q++)
#pragma unroll
        for (int i = 0; i < 4; i++) acc[q][i] = 0;
#pragma unroll 1
    for (uint32_t it = 0; it < iters; it++) {
#pragma unroll
        for (int q = 0; q < 8; q++) mma_u8(acc[q], a, b0 + q, b1);
    }
    int s = 0;
#pragma unroll
    for (int q = 0; q < 8; q++)
#pragma unroll
        for (int i = 0; i < 4; i++) s ^= acc[q][i];
    if (s == 0x12345678) out[0] = s;
}

// ---- per-warp shared-memory areas ---------------------------------------------------------------------------
constexpr int kARowStride = 48;                 // bytes per sponge row of an A tile (32 used; 48 avoids ldmatrix conflicts)
constexpr int kATile = 32 * kARowStride;        // one value of 32 sponges
constexpr int kCRowStride = 80;                 // bytes per sponge row of the column-sum exchange (64 used)
constexpr int kCTile = 32 * kCRowStride;
constexpr int kWarpSmem = kATile + kCTile;      // 4096 bytes per warp
constexpr int kTabBytes = 5 * 4 * 256;          // 5 tables (M00, w1, w2, v1, v2) x 4 n-tiles x (32 lanes x 8 bytes)

// x (8 words, this thread's sponge) -> A fragments of the two 16-sponge m-tiles of the warp
__device__ __forceinline__ void pack_a(uint32_t (&a)[2][4], const uint32_t* x, unsigned char* a_tile, int lane) {
    uint4* row = reinterpret_cast<uint4*>(a_tile + lane * kARowStride);
    row[0] = make_uint4(x[0], x[1], x[2], x[3]);
    row[1] = make_uint4(x[4], x[5], x[6], x[7]);
    __syncwarp();
    // matrix q of the x4 load: rows 8*(q&1).. of the m-tile, bytes 16*(q>>1)..; lane L supplies row (L&7) of matrix L>>3
    const int q = lane >> 3, r = lane & 7;
#pragma unroll
    for (int mt = 0; mt < 2; mt++) {
        const unsigned char* src = a_tile + (mt * 16 + (q & 1) * 8 + r) * kARowStride + (q >> 1) * 16;
        ldmatrix_x4(a[mt], (uint32_t)__cvta_generic_to_shared(src));
    }
    __syncwarp();
}

// column sums of one output value (2 m-tiles x 4 n-tiles x 4) -> 8 x 32-bit limbs of THIS thread's sponge:
// pair-combine, exchange through shared memory, carry sweep, one 32-bit Montgomery row
__device__ __forceinline__ void sweep(uint32_t* out, const int (&acc)[2][4][4], unsigned char* c_tile, int lane) {
    const int g = lane >> 2, tg = lane & 3;
#pragma unroll
    for (int mt = 0; mt < 2; mt++)
#pragma unroll
        for (int h = 0; h < 2; h++) {
            // row g + 8h + 16mt, columns 8nt + 2tg (+1): P[nt] = c_even + 256 c_odd
            uint32_t P[4];
#pragma unroll
            for (int nt = 0; nt < 4; nt++) P[nt] = (uint32_t)acc[mt][nt][2 * h] + ((uint32_t)acc[mt][nt][2 * h + 1] << 8);
            *reinterpret_cast<uint4*>(c_tile + (g + 8 * h + 16 * mt) * kCRowStride + tg * 16) = make_uint4(P[0], P[1], P[2], P[3]);
        }
    __syncwarp();
    uint32_t Q[4][4];   // [tg][nt] of this thread's sponge (row = lane)
#pragma unroll
    for (int k = 0; k < 4; k++) {
        const uint4 v = *reinterpret_cast<const uint4*>(c_tile + lane * kCRowStride + k * 16);
        Q[k][0] = v.x; Q[k][1] = v.y; Q[k][2] = v.z; Q[k][3] = v.w;
    }
    __syncwarp();
    // value = sum Q[tg][nt] 2^(64 nt + 16 tg)
    uint32_t v[9];
    unsigned long long carry = 0;
#pragma unroll
    for (int nt = 0; nt < 4; nt++) {
        unsigned __int128 w = (unsigned __int128)Q[0][nt] + ((unsigned __int128)Q[1][nt] << 16) + ((unsigned __int128)Q[2][nt] << 32) +
                              ((unsigned __int128)Q[3][nt] << 48) + carry;
        v[2 * nt] = (uint32_t)w;
        v[2 * nt + 1] = (uint32_t)(w >> 32);
        carry = (unsigned long long)(w >> 64);
    }
    v[8] = (uint32_t)carry;
    // one Montgomery row: (v + m p) / 2^32
    const uint32_t m = v[0] * F::NP0;
    uint32_t lo[8] = {v[0], v[1], v[2], v[3], v[4], v[5], v[6], v[7]};
    uint32_t hi[8] = {v[1], v[2], v[3], v[4], v[5], v[6], v[7], v[8]};
    uint32_t c8 = 0, top = 0;
    mad4(lo, F::P0, F::P2, F::P4, F::P6, m, c8);     // columns (0,1)..(6,7)
    mad4(hi, F::P1, F::P3, F::P5, F::P7, m, top);    // columns (1,2)..(7,8)
    // result limbs = columns 1..8 of lo-aligned + hi-aligned parts (mock: keep the dependency, merge once)
    uint32_t z[8] = {lo[0], lo[1], lo[2], lo[3], lo[4], lo[5], lo[6], lo[7]};
    z[0] = 0;
    merge8(out, hi, z);
    out[7] += c8 + top;
}

template <int MODE>
__global__ void __launch_bounds__(512, 1) k_round(uint32_t* out, const unsigned char* tab_g, uint32_t iters) {
    extern __shared__ __align__(16) unsigned char smem[];
    unsigned char* tab = smem;                                    // kTabBytes, shared by the block
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    unsigned char* a_tile = smem + kTabBytes + warp * kWarpSmem;
    unsigned char* c_tile = a_tile + kATile;
    for (int i = threadIdx.x; i < kTabBytes / 16; i += blockDim.x)
        reinterpret_cast<uint4*>(tab)[i] = reinterpret_cast<const uint4*>(tab_g)[i];
    __syncthreads();

    uint32_t x[8], l1[8], l2[8];
#pragma unroll
    for (int k = 0; k < 8; k++) {
        x[k] = (threadIdx.x + 1) * 2654435761u + k * 40503u + blockIdx.x;
        l1[k] = x[k] * 3u + 1u;
        l2[k] = x[k] * 5u + 7u;
    }
    x[7] &= 0x0fffffffu;
    l1[7] &= 0x0fffffffu;
    l2[7] &= 0x0fffffffu;
    uint32_t aL1[2][4], aL2[2][4];
    pack_a(aL1, l1, a_tile, lane);
    pack_a(aL2, l2, a_tile, lane);

#pragma unroll 1
    for (uint32_t it = 0; it < iters; it++) {
        if (MODE != 2) {
            uint32_t y[8];
            fp_sqr<F, false>(y, x);
            fp_sqr<F, false>(y, y);
            mont_mul<F, false>(x, y, x);
        }
        if (MODE != 0) {
            uint32_t aY[2][4];
            pack_a(aY, x, a_tile, lane);
            // row 0 = M00 y + w1 lane1 + w2 lane2  (tables 0, 1, 2)
            int acc[2][4][4];
#pragma unroll
            for (int mt = 0; mt < 2; mt++)
#pragma unroll
                for (int nt = 0; nt < 4; nt++)
#pragma unroll
                    for (int i = 0; i < 4; i++) acc[mt][nt][i] = 0;
#pragma unroll
            for (int ks = 0; ks < 3; ks++)
#pragma unroll
                for (int nt = 0; nt < 4; nt++) {
                    const uint2 b = *reinterpret_cast<const uint2*>(tab + (ks * 4 + nt) * 256 + lane * 8);
                    mma_u8(acc[0][nt], ks == 0 ? aY[0] : ks == 1 ? aL1[0] : aL2[0], b.x, b.y);
                    mma_u8(acc[1][nt], ks == 0 ? aY[1] : ks == 1 ? aL1[1] : aL2[1], b.x, b.y);
                }
            uint32_t nx[8];
            sweep(nx, acc, c_tile, lane);
            // lane updates: lane_j += v_j y  (tables 3, 4), then re-pack the lanes as A fragments
#pragma unroll
            for (int j = 0; j < 2; j++) {
#pragma unroll
                for (int mt = 0; mt < 2; mt++)
#pragma unroll
                    for (int nt = 0; nt < 4; nt++)
#pragma unroll
                        for (int i = 0; i < 4; i++) acc[mt][nt][i] = 0;
#pragma unroll
                for (int nt = 0; nt < 4; nt++) {
                    const uint2 b = *reinterpret_cast<const uint2*>(tab + ((3 + j) * 4 + nt) * 256 + lane * 8);
                    mma_u8(acc[0][nt], aY[0], b.x, b.y);
                    mma_u8(acc[1][nt], aY[1], b.x, b.y);
                }
                uint32_t d[8];
                sweep(d, acc, c_tile, lane);
                uint32_t* l = j == 0 ? l1 : l2;
                add8(l, l, d);
                l[7] &= 0x0fffffffu;
                if (j == 0) pack_a(aL1, l, a_tile, lane);
                else pack_a(aL2, l, a_tile, lane);
            }
#pragma unroll
            for (int k = 0; k < 8; k++) x[k] = nx[k];
            x[7] &= 0x0fffffffu;   // keep the mock value inside the S-box's operand range
        }
    }
    uint32_t s = 0;
#pragma unroll
    for (int k = 0; k < 8; k++) s ^= x[k] ^ l1[k] ^ l2[k];
    if (s == 0x12345678u) out[0] = s;
}

static double clock_mhz() {
    int khz = 0;
    cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, 0);
    return khz / 1000.0;
}

template <class L>
static double time_ms(L launch, int reps) {
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    launch();
    cudaDeviceSynchronize();
    cudaEventRecord(e0);
    for (int r = 0; r < reps; r++) launch();
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms = 0;
    cudaEventElapsedTime(&ms, e0, e1);
    return ms / reps;
}

int main() {
    cudaDeviceProp prop;
    cudaGetDeviceProperties(&prop, 0);
    const int sms = prop.multiProcessorCount;
    const double mhz = clock_mhz();
    printf("%s: %d SMs, %.0f MHz (nominal; cycles below use it)\n", prop.name, sms, mhz);
    uint32_t* d_out;
    cudaMalloc(&d_out, 4096);
    std::vector<unsigned char> tab(kTabBytes);
    for (size_t i = 0; i < tab.size(); i++) tab[i] = (unsigned char)(i * 197u + 13u);
    unsigned char* d_tab;
    cudaMalloc(&d_tab, kTabBytes);
    cudaMemcpy(d_tab, tab.data(), kTabBytes, cudaMemcpyHostToDevice);
    const size_t smem = kTabBytes + 16 * kWarpSmem;
    cudaFuncSetAttribute(k_round<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    cudaFuncSetAttribute(k_round<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    cudaFuncSetAttribute(k_round<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);

    {
        const uint32_t iters = 4096;
        const double ms = time_ms([&] { k_imma<<<sms, 512>>>((int*)d_out, iters); }, 3);
        const double cyc = ms * 1e-3 * mhz * 1e6;
        // 16 warps per SM = 4 per sub-partition, 8 mma per iteration each
        printf("IMMA m16n8k32 u8: %.2f cycles per mma per sub-partition (%.1f dense int8 TOP/s on the GPU)\n",
               cyc / (iters * 8.0 * 4.0), sms * 4.0 * 8192.0 * 2.0 / (cyc / (iters * 8.0 * 4.0)) * mhz * 1e6 / 1e12 / 2.0);
    }
    const uint32_t iters = 512;
    const char* names[3] = {"S-box x^5 alone (integer, lazy)", "mock partial round t=3: S-box + tensor-core linear layer",
                            "tensor-core linear layer alone (pack, 40 mma, 3 sweeps, lane re-pack)"};
    for (int mode = 0; mode < 3; mode++) {
        double ms;
        if (mode == 0) ms = time_ms([&] { k_round<0><<<sms, 512, smem>>>(d_out, d_tab, iters); }, 3);
        else if (mode == 1) ms = time_ms([&] { k_round<1><<<sms, 512, smem>>>(d_out, d_tab, iters); }, 3);
        else ms = time_ms([&] { k_round<2><<<sms, 512, smem>>>(d_out, d_tab, iters); }, 3);
        const double cyc = ms * 1e-3 * mhz * 1e6;
        printf("%-75s %8.0f cycles per warp-round per sub-partition  (%.3f ms)\n", names[mode], cyc / (iters * 4.0), ms);
    }
    cudaError_t e = cudaDeviceSynchronize();
    printf("status: %s\n", cudaGetErrorString(e));
    return e == cudaSuccess ? 0 : 1;
}
