// VARIANT 6: the optimised Poseidon permutation with the S-box on the integer multiplier and the whole linear
// layer on the TENSOR CORES.  Same function, bit for bit, as poseidon/mod.rs:95-118.
//
// A t x t MDS product over 254-bit integers is not a contraction -- but a product by a ROUND CONSTANT is, once
// the constant is expanded into a digit table: for a constant c the host stores the 32 x 32 byte matrix
//     Tab_c[k][col] = byte `col` of ( c * 256^k * 2^32 mod p ),        k, col = 0..31
// and then, for any 256-bit x with bytes x_k,
//     x * c * 2^32  ==  sum_col 256^col * ( sum_k x_k * Tab_c[k][col] )      (mod p).
// Over the 32 sponges of a warp the inner sums are an integer GEMM  [32 sponges x 32 T bytes] x [32 T x 32]  with
// u8 operands and s32 accumulation (column sums < 32 T 255^2 < 2^24 for T <= 5): mma.sync.m16n8k32.u8.u8.s32,
// two m-tiles per warp, A fragments built from the thread-per-sponge registers through shared memory + ldmatrix,
// B fragments (the tables, pre-arranged by the host in fragment order) read from global memory / L1 with one
// 64-bit load per pair of MMAs.  What remains per output value is a "sweep" by the thread that owns the sponge:
// the 32 byte-weighted column sums come back through shared memory, are folded into nine 32-bit limbs (value
// < 2^272.1) and ONE 32-bit Montgomery row divides the 2^32 back out: result < p + 2^241, congruent to the exact
// field value -- a valid input for the lazy S-box (fp256.cuh) and for the next round's byte expansion.
// Additive round constants enter as the initial values of the accumulators (bytes of K * 2^32 mod p).
//
// Round structure: the regrouped (sparse) form of opt_constants.cpp, as in variants 1-5.  In the partial
// rounds lane j >= 1 is updated as  s_j' = 1 * s_j + v_j * y  with BOTH terms through the tensor cores (the
// identity's table costs one more k-step and returns a value that is bounded again, so lanes never grow), and
// lives only as A fragments between rounds.
//
// Nothing here touches the FP64 pipe and the integer multiplier sees only the S-box and 8 multiplies per
// sweep: per partial round (t = 3) 352 + 24 IMAD.WIDE against 730 in variant 3, and about 900 instructions
// against 1 528 in variant 5.
#pragma once
#include "poseidon_core.cuh"

#ifndef PSD_TC_PIPELINE
#define PSD_TC_PIPELINE 1
#endif

namespace psd {

constexpr uint32_t kTcFragBytes = 1024;   // one constant: 4 n-tiles x 32 lanes x (b0, b1)
constexpr uint32_t kTcKBytes = 256;       // one additive constant: 4 n-tiles x 4 column pairs x (k0, k1, k0, k1) s32
constexpr uint32_t kTc5KBytes = 32;       // layout 1: one additive constant = the eight 32-bit limbs of K 2^32 mod p
// per-warp shared-memory scratch: two A tiles (32 sponges x 32 bytes, rows padded to 48) and two column-sum
// exchange tiles (32 sponges x 16 words, rows padded to 20 words); consecutive uses alternate between the two
// copies so that only the write -> read direction needs a warp barrier
constexpr uint32_t kTcARow = 48, kTcATile = 32 * kTcARow;
constexpr uint32_t kTcCRow = 80, kTcCTile = 32 * kTcCRow;
constexpr uint32_t kTcWarpScratch = 2 * kTcATile + 2 * kTcCTile;   // 8 192 bytes

// Where variant 6 finds its tables (byte offsets into the global image `tab`), see tc_tables.cpp.
struct TcDesc {
    uint32_t ok;
    uint32_t layout;      // 0: mma.sync fragment order (variant 6); 1: tcgen05 K-major core-matrix blocks, 64-row tables for the unreduced
                          // S-box values, tables that share an operand stacked along N (variant 7)
    uint32_t mat_off;     // [T][T] fragments of the dense matrix ([0][0] doubles as M00)
    uint32_t pre_off;     // [T][T] fragments of M_pre
    uint32_t ident_off;   // fragment of the constant 1
    uint32_t part_off;    // [partial][2(T-1)] fragments: w_1.., v_1.. per partial round
    uint32_t pair_off;    // [partial / 2] fragments: c_q = sum_j w_j(i+1) v_j(i) for the q-th PAIR of partial rounds (i, i+1)
    uint32_t k_off;       // [rounds][T] additive constants (accumulator start values), execution order
    // layout 1 only (poseidon_tc5.cuh has the table formats)
    uint32_t red_off;     // 64-row table of the constant 1: the reduction of an unreduced square
    uint32_t m00_off;     // 64-row table of M00 on its own
};

// the lazy S-box of fp256.cuh / poseidon_core.cuh on a value < p + 2^241 (BN254), else the reducing one after one
// conditional subtraction
template <class F, int ALPHA>
PSD_DEV void tc_sbox(uint32_t* x) {
    constexpr bool kLazy = (F::kBits == 254);
    uint32_t in[8];
    if (kLazy) {
#pragma unroll
        for (int k = 0; k < 8; k++) in[k] = x[k];
    } else {
        cond_sub_p<F>(in, x);
    }
    uint32_t y[8];
    fp_sqr<F, !kLazy>(y, in);
#pragma unroll
    for (int q = 1; q < (ALPHA == 5 ? 2 : 4); q++) fp_sqr<F, !kLazy>(y, y);
    mont_mul<F, !kLazy>(x, y, in);
}

#if defined(__CUDACC__)

struct TcWarp {
    unsigned char* scratch;   // this warp's kTcWarpScratch bytes of shared memory
    const unsigned char* tab; // the context's table image in global memory
    int lane;
    uint32_t a_use, c_use;    // which copy of the A / C tile the next use takes
};

__device__ __forceinline__ void tc_mma(int (&d)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
    asm("mma.sync.aligned.m16n8k32.row.col.s32.u8.u8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
        : "+r"(d[0]), "+r"(d[1]), "+r"(d[2]), "+r"(d[3])
        : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}

// x (8 words = 32 little-endian bytes of this thread's sponge) -> A fragments of the warp's two 16-sponge m-tiles
__device__ __forceinline__ void tc_pack(uint32_t (&a)[2][4], const uint32_t* x, TcWarp& w) {
    unsigned char* tile = w.scratch + (w.a_use & 1) * kTcATile;
    w.a_use++;
    uint4* row = reinterpret_cast<uint4*>(tile + w.lane * kTcARow);
    row[0] = make_uint4(x[0], x[1], x[2], x[3]);
    row[1] = make_uint4(x[4], x[5], x[6], x[7]);
    __syncwarp();
    // ldmatrix.x4: lane L supplies the address of row (L & 7) of 8x8 b16 matrix (L >> 3); matrix q covers sponges
    // 8 (q & 1) .. + 7 of the m-tile and bytes 16 (q >> 1) .. + 15, which is the a0..a3 order of m16n8k32
    const int q = w.lane >> 3, r = w.lane & 7;
#pragma unroll
    for (int mt = 0; mt < 2; mt++) {
        const unsigned char* src = tile + (mt * 16 + (q & 1) * 8 + r) * kTcARow + (q >> 1) * 16;
        const uint32_t addr = (uint32_t)__cvta_generic_to_shared(src);
        asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
                     : "=r"(a[mt][0]), "=r"(a[mt][1]), "=r"(a[mt][2]), "=r"(a[mt][3])
                     : "r"(addr)
                     : "memory");
    }
}

// d = a x b + c with c given separately (the first k-step of an output value: c = the additive constant's bytes,
// so the accumulators need no initialising moves; the table stores (k0, k1, k0, k1) so that one 128-bit load
// delivers the four consecutive registers the instruction wants)
__device__ __forceinline__ void tc_mma_c(int (&d)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1, const int4& c) {
    asm("mma.sync.aligned.m16n8k32.row.col.s32.u8.u8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%10,%11,%12,%13};"
        : "=r"(d[0]), "=r"(d[1]), "=r"(d[2]), "=r"(d[3])
        : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1), "r"(c.x), "r"(c.y), "r"(c.z), "r"(c.w));
}

// accumulators = K + bytes(value behind the A fragments) x Tab_c: the additive constant's bytes (columns 8 nt + 2 tg,
// + 1 of every row this thread holds; the same for all its rows) enter as the C operand of the first k-step
__device__ __forceinline__ void tc_mac_first(int (&acc)[2][4][4], const uint32_t (&a)[2][4], const unsigned char* frag,
                                             const unsigned char* kfrag, const TcWarp& w) {
    const int tg = w.lane & 3;
#pragma unroll
    for (int nt = 0; nt < 4; nt++) {
        const int4 k = __ldg(reinterpret_cast<const int4*>(kfrag) + nt * 4 + tg);
        const uint2 b = __ldg(reinterpret_cast<const uint2*>(frag) + nt * 32 + w.lane);
        tc_mma_c(acc[0][nt], a[0], b.x, b.y, k);
        tc_mma_c(acc[1][nt], a[1], b.x, b.y, k);
    }
}

// accumulators += bytes(value behind the A fragments) x Tab_c, c given by its fragment block
__device__ __forceinline__ void tc_mac(int (&acc)[2][4][4], const uint32_t (&a)[2][4], const unsigned char* frag, const TcWarp& w) {
#pragma unroll
    for (int nt = 0; nt < 4; nt++) {
        const uint2 b = __ldg(reinterpret_cast<const uint2*>(frag) + nt * 32 + w.lane);
        tc_mma(acc[0][nt], a[0], b.x, b.y);
        tc_mma(acc[1][nt], a[1], b.x, b.y);
    }
}

// column sums of one output value -> 8 x 32-bit limbs of THIS thread's sponge, value < p + 2^241:
// pair-combine, exchange through shared memory, carry sweep, one 32-bit Montgomery row
template <class F>
__device__ __forceinline__ void tc_sweep(uint32_t* out, const int (&acc)[2][4][4], TcWarp& w) {
    unsigned char* tile = w.scratch + 2 * kTcATile + (w.c_use & 1) * kTcCTile;
    w.c_use++;
    const int g = w.lane >> 2, tg = w.lane & 3;
#pragma unroll
    for (int mt = 0; mt < 2; mt++)
#pragma unroll
        for (int h = 0; h < 2; h++) {
            // sponge g + 8 h + 16 mt, columns 8 nt + 2 tg and + 1:  P = c_even + 256 c_odd  (< 2^32: sums < 2^32 / 257)
            uint32_t P[4];
#pragma unroll
            for (int nt = 0; nt < 4; nt++) P[nt] = (uint32_t)acc[mt][nt][2 * h] + ((uint32_t)acc[mt][nt][2 * h + 1] << 8);
            *reinterpret_cast<uint4*>(tile + (g + 8 * h + 16 * mt) * kTcCRow + tg * 16) = make_uint4(P[0], P[1], P[2], P[3]);
        }
    __syncwarp();
    uint32_t Q[4][4];   // [column pair tg][n-tile] of this thread's sponge
#pragma unroll
    for (int k = 0; k < 4; k++) {
        const uint4 v = *reinterpret_cast<const uint4*>(tile + w.lane * kTcCRow + k * 16);
        Q[k][0] = v.x; Q[k][1] = v.y; Q[k][2] = v.z; Q[k][3] = v.w;
    }
    // value = sum_{nt, k} Q[k][nt] 2^(64 nt + 16 k) = X + (Y << 16) with
    //   X = limbs (Q[0][0], Q[2][0], Q[0][1], Q[2][1], ...)   (column pairs at bit offsets 0 and 32 of each 64-bit word)
    //   Y = limbs (Q[1][0], Q[3][0], Q[1][1], Q[3][1], ...)   (offsets 16 and 48)
    // every Q is below 2^32, so X and Y are plain 256-bit numbers: nine funnel shifts and one carry chain
    uint32_t v[9];
    {
        uint32_t ys[9];
        ys[0] = Q[1][0] << 16;
#pragma unroll
        for (int k = 1; k < 8; k++) {
            const uint32_t lo = (k & 1) ? Q[1][k >> 1] : Q[3][(k >> 1) - 1];   // limb k - 1 of Y
            const uint32_t hi = (k & 1) ? Q[3][k >> 1] : Q[1][k >> 1];         // limb k of Y
            ys[k] = __funnelshift_l(lo, hi, 16);
        }
        ys[8] = Q[3][3] >> 16;
        asm("add.cc.u32 %0, %9, %18;\n\t"
            "addc.cc.u32 %1, %10, %19;\n\t"
            "addc.cc.u32 %2, %11, %20;\n\t"
            "addc.cc.u32 %3, %12, %21;\n\t"
            "addc.cc.u32 %4, %13, %22;\n\t"
            "addc.cc.u32 %5, %14, %23;\n\t"
            "addc.cc.u32 %6, %15, %24;\n\t"
            "addc.cc.u32 %7, %16, %25;\n\t"
            "addc.u32 %8, %17, 0;"
            : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8])
            : "r"(Q[0][0]), "r"(Q[2][0]), "r"(Q[0][1]), "r"(Q[2][1]), "r"(Q[0][2]), "r"(Q[2][2]), "r"(Q[0][3]), "r"(Q[2][3]),
              "r"(ys[8]), "r"(ys[0]), "r"(ys[1]), "r"(ys[2]), "r"(ys[3]), "r"(ys[4]), "r"(ys[5]), "r"(ys[6]), "r"(ys[7]));
    }
    // Montgomery row: (V + m p) / 2^32, m = -V / p mod 2^32.  First the even limbs of p onto columns 0..7 (column 0
    // becomes zero), then the odd limbs onto columns 1..8; the carry out of column 8 is provably zero
    // (V + m p < 2^273 + 2^286 < 2^288).
    const uint32_t m = v[0] * np0<F>();
    uint32_t c8 = 0;
    mad4(v, plimb<F, 0>(), plimb<F, 2>(), plimb<F, 4>(), plimb<F, 6>(), m, c8);
    uint32_t x[8] = {v[1], v[2], v[3], v[4], v[5], v[6], v[7], v[8] + c8};
    mad4_nc(x, plimb<F, 1>(), plimb<F, 3>(), plimb<F, 5>(), plimb<F, 7>(), m);
#pragma unroll
    for (int k = 0; k < 8; k++) out[k] = x[k];
}

// One full round on limb states: T S-boxes, then T rows of `mat` (each: constant start values, T k-steps, sweep).
// Free functions rather than lambdas: a lambda's call operator is not force-inlined, and an outlined round
// would pass its arrays through local memory.
template <class F, int T, int ALPHA>
__device__ __forceinline__ void tc_full_round(uint32_t (&st)[T][8], const unsigned char* mat, const unsigned char* kp, TcWarp& w) {
    uint32_t aX[T][2][4];
#pragma unroll
    for (int j = 0; j < T; j++) {
        tc_sbox<F, ALPHA>(st[j]);
        tc_pack(aX[j], st[j], w);
    }
    int acc[2][4][4];
#pragma unroll
    for (int i = 0; i < T; i++) {
        tc_mac_first(acc, aX[0], mat + (i * T) * kTcFragBytes, kp + i * kTcKBytes, w);
#pragma unroll
        for (int j = 1; j < T; j++) tc_mac(acc, aX[j], mat + (i * T + j) * kTcFragBytes, w);
        tc_sweep<F>(st[i], acc, w);
    }
}

// Partial rounds, software-pipelined.  Only lane 0's chain  S-box -> row 0 -> S-box  is serial; the lane updates of
// round i depend on y_i alone, so they are written NEXT TO the S-box of round i + 1: the S-box keeps the integer
// multiplier busy while the lane updates use the tensor cores and shared memory.
//   tc_row0:   st0 <- M00 y + sum_j w_j s_j + k        (y, lanes given as A fragments)
//   tc_lanes:  s_j <- 1 s_j + v_j y + k_j               (REPACK: result back to A fragments, else to limbs)
template <class F, int T>
__device__ __forceinline__ void tc_row0(uint32_t* st0, const uint32_t (&aY)[2][4], const uint32_t (&aL)[T - 1][2][4],
                                        const unsigned char* m00, const unsigned char* part, const unsigned char* kp, TcWarp& w) {
    int acc[2][4][4];
    tc_mac_first(acc, aY, m00, kp, w);
#pragma unroll
    for (int j = 1; j < T; j++) tc_mac(acc, aL[j - 1], part + (j - 1) * kTcFragBytes, w);
    tc_sweep<F>(st0, acc, w);
}

template <class F, int T, bool REPACK>
__device__ __forceinline__ void tc_lanes(uint32_t (&st)[T][8], const uint32_t (&aY)[2][4], uint32_t (&aL)[T - 1][2][4],
                                         const unsigned char* ident, const unsigned char* part, const unsigned char* kp,
                                         TcWarp& w) {
#pragma unroll
    for (int j = 1; j < T; j++) {
        int acc[2][4][4];
        tc_mac_first(acc, aY, part + (T - 1 + j - 1) * kTcFragBytes, kp + j * kTcKBytes, w);
        tc_mac(acc, aL[j - 1], ident, w);
        uint32_t lane[8];
        tc_sweep<F>(lane, acc, w);
        if (REPACK) {
            tc_pack(aL[j - 1], lane, w);
        } else {
#pragma unroll
            for (int k = 0; k < 8; k++) st[j][k] = lane[k];
        }
    }
}

// Partial rounds are run in PAIRS (i, i+1): between partial rounds the lanes change only by a multiple of the S-box
// output, so the second row 0 of a pair can take the lanes as they were BEFORE the pair plus one extra term,
//   row0(i+1) = M00 y1 + (sum_j w_j(i+1) v_j(i)) y0 + sum_j w_j(i+1) s_j + k,
// and the lanes are brought up to date once per pair:  s_j <- 1 s_j + v_j(i) y0 + v_j(i+1) y1.  Per pair and lane that is
// one sweep and one re-pack instead of two, for one more k-step in the second row.
template <class F, int T>
__device__ __forceinline__ void tc_row0_second(uint32_t* st0, const uint32_t (&aY1)[2][4], const uint32_t (&aY0)[2][4],
                                               const uint32_t (&aL)[T - 1][2][4], const unsigned char* m00, const unsigned char* pair,
                                               const unsigned char* part_next, const unsigned char* kp_next, TcWarp& w) {
    int acc[2][4][4];
    tc_mac_first(acc, aY1, m00, kp_next, w);
    tc_mac(acc, aY0, pair, w);
#pragma unroll
    for (int j = 1; j < T; j++) tc_mac(acc, aL[j - 1], part_next + (j - 1) * kTcFragBytes, w);
    tc_sweep<F>(st0, acc, w);
}

template <class F, int T, bool REPACK>
__device__ __forceinline__ void tc_lanes_pair(uint32_t (&st)[T][8], const uint32_t (&aY0)[2][4], const uint32_t (&aY1)[2][4],
                                              uint32_t (&aL)[T - 1][2][4], const unsigned char* ident, const unsigned char* part,
                                              const unsigned char* part_next, const unsigned char* kp_next, TcWarp& w) {
#pragma unroll
    for (int j = 1; j < T; j++) {
        int acc[2][4][4];
        tc_mac_first(acc, aY1, part_next + (T - 1 + j - 1) * kTcFragBytes, kp_next + j * kTcKBytes, w);
        tc_mac(acc, aY0, part + (T - 1 + j - 1) * kTcFragBytes, w);
        tc_mac(acc, aL[j - 1], ident, w);
        uint32_t lane[8];
        tc_sweep<F>(lane, acc, w);
        if (REPACK) {
            tc_pack(aL[j - 1], lane, w);
        } else {
#pragma unroll
            for (int k = 0; k < 8; k++) st[j][k] = lane[k];
        }
    }
}

template <class F, int T, int ALPHA>
__device__ __forceinline__ void permute_tc(uint32_t (&st)[T][8], const PermDesc& d, const TcDesc& td, TcWarp& w) {
    static_assert(ALPHA == 5 || ALPHA == 17, "variant 6: alpha 5 or 17");
    // round-0 constants on the integer side
#pragma unroll
    for (int j = 0; j < T; j++) {
        uint32_t c[8];
        cload8(c, d.ark_off + j * 8);
        fp_add<F>(st[j], st[j], c);
    }
    const uint32_t half = d.half_full, rp = d.partial;
    const unsigned char* kp = w.tab + td.k_off;

#pragma unroll 1
    for (uint32_t r = 0; r < half; r++) {
        tc_full_round<F, T, ALPHA>(st, w.tab + ((r + 1 == half && rp) ? td.pre_off : td.mat_off), kp, w);
        kp += T * kTcKBytes;
    }

    if (rp) {
        constexpr uint32_t kRoundK = T * kTcKBytes, kRoundPart = 2 * (T - 1) * kTcFragBytes;
        uint32_t aL[T - 1][2][4], aY0[2][4];
#pragma unroll
        for (int j = 1; j < T; j++) tc_pack(aL[j - 1], st[j], w);
        const unsigned char* part = w.tab + td.part_off;
        const unsigned char* pair = w.tab + td.pair_off;
        const unsigned char* m00 = w.tab + td.mat_off;
        const unsigned char* ident = w.tab + td.ident_off;
        const uint32_t pairs = rp >> 1;
        tc_sbox<F, ALPHA>(st[0]);
        tc_pack(aY0, st[0], w);
        if (rp & 1) {
            // an odd number of partial rounds: the first one on its own
            tc_row0<F, T>(st[0], aY0, aL, m00, part, kp, w);
            if (pairs) {
                tc_sbox<F, ALPHA>(st[0]);
                tc_lanes<F, T, true>(st, aY0, aL, ident, part, kp, w);
                tc_pack(aY0, st[0], w);
            } else {
                tc_lanes<F, T, false>(st, aY0, aL, ident, part, kp, w);
            }
            kp += kRoundK;
            part += kRoundPart;
        }
        if (pairs) {
            uint32_t aY1[2][4];
#pragma unroll 1
            for (uint32_t q = 0; q + 1 < pairs; q++) {
                tc_row0<F, T>(st[0], aY0, aL, m00, part, kp, w);
                tc_sbox<F, ALPHA>(st[0]);
                tc_pack(aY1, st[0], w);
                tc_row0_second<F, T>(st[0], aY1, aY0, aL, m00, pair, part + kRoundPart, kp + kRoundK, w);
                // independent of each other: the S-box of the next pair's first round and this pair's lane updates
                tc_sbox<F, ALPHA>(st[0]);
                tc_lanes_pair<F, T, true>(st, aY0, aY1, aL, ident, part, part + kRoundPart, kp + kRoundK, w);
                tc_pack(aY0, st[0], w);
                kp += 2 * kRoundK;
                part += 2 * kRoundPart;
                pair += kTcFragBytes;
            }
            // last pair: the updated lanes are needed as limbs again (full rounds follow)
            tc_row0<F, T>(st[0], aY0, aL, m00, part, kp, w);
            tc_sbox<F, ALPHA>(st[0]);
            tc_pack(aY1, st[0], w);
            tc_row0_second<F, T>(st[0], aY1, aY0, aL, m00, pair, part + kRoundPart, kp + kRoundK, w);
            tc_lanes_pair<F, T, false>(st, aY0, aY1, aL, ident, part, part + kRoundPart, kp + kRoundK, w);
            kp += 2 * kRoundK;
        }
    }

#pragma unroll 1
    for (uint32_t r = half + rp; r < d.total; r++) {
        tc_full_round<F, T, ALPHA>(st, w.tab + td.mat_off, kp, w);
        kp += T * kTcKBytes;
    }

    // sweeps leave values < p + 2^241: one conditional subtraction gives the canonical residue
#pragma unroll
    for (int j = 0; j < T; j++) {
        uint32_t x[8];
#pragma unroll
        for (int k = 0; k < 8; k++) x[k] = st[j][k];
        cond_sub_p<F>(st[j], x);
    }
}

#endif  // __CUDACC__

}  // namespace psd
