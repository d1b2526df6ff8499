// VARIANT 7: variant 6's contraction (products by round constants as u8 x u8 -> s32 matrix products against byte-digit
// tables, see poseidon_tc.cuh) on the 5th-generation tensor cores: tcgen05.mma.kind::i8 with both operands in shared
// memory and the accumulators in TMEM.
//
// Why: with mma.sync every warp issues its own MMAs (56 per partial round, 8.8 issue cycles each), builds A fragments with
// ldmatrix, loads B fragments, and brings the column sums back to the owning thread through shared memory.  Here a thread
// block is four TILES of 128 sponges (4 warps each); per tile
//   * every thread writes the 32 bytes of a value ONCE into the tile's A slot (K-major core-matrix layout: 16-byte chunk c of
//     row r at c * 2048 + r * 16),
//   * ONE thread issues the MMAs of a whole step (M = 128 sponges, N = 32 columns, K = 32 bytes per instruction; a step is
//     3..10 instructions) and commits them to the tile's mbarrier,
//   * the accumulators live in the tile's 128 TMEM columns, and tcgen05.ld.32x32b hands thread r the 32 column sums of ITS
//     OWN row -- no exchange, no fragments,
//   * the round's tables (1 KiB per constant, the B operand as stored by the host) are staged per tile with cp.async into
//     a double buffer one pair of partial rounds ahead; the dense matrices stay in shared memory for the whole block.
// The arithmetic -- tables, column sums, carry sweep, one 32-bit Montgomery row, lazy S-box, pairs of partial rounds -- is
// exactly variant 6's, so the two produce the same intermediate values and the host model of tests/emu covers both.
#pragma once
#include "poseidon_tc.cuh"

namespace psd {

constexpr uint32_t kTc5ASlot = 4096;    // one value of a 128-sponge tile: [16-byte chunk 0..1][row 0..127][16 bytes]
constexpr uint32_t kTc5Tab = 1024;      // one table: [k / 16][column 0..31][k % 16]
constexpr uint32_t kTc5TileRows = 128;

// shared-memory bytes of the variant-7 kernel behind the staged constant image, for T lanes and 4 tiles
#if defined(__CUDACC__)
#define PSD_HD __host__ __device__
#else
#define PSD_HD
#endif
PSD_HD constexpr uint32_t tc5_static_bytes(uint32_t t) { return (2 * t * t + 1) * kTc5Tab; }
PSD_HD constexpr uint32_t tc5_pair_tables(uint32_t t) { return 4 * (t - 1) + 1; }
PSD_HD constexpr uint32_t tc5_tile_bytes(uint32_t t) { return (t + 1) * kTc5ASlot + 2 * tc5_pair_tables(t) * kTc5Tab; }
PSD_HD constexpr uint32_t tc5_smem_bytes(uint32_t t, uint32_t tiles) { return tc5_static_bytes(t) + tiles * tc5_tile_bytes(t) + 128; }

#if defined(__CUDACC__)

struct Tc5Tile {
    uint32_t tmem;            // TMEM address of this tile's column 0 (lane field 0)
    uint32_t a_u32, tab_u32;  // shared-memory addresses of the tile's A slots / table double buffer
    uint32_t static_u32;      // ... of the block's static tables: mat [T*T], pre [T*T], ident
    // Two mbarriers per tile: a step commits twice (row 0 first, the lanes behind it), and a waiter may be at most one
    // phase behind an mbarrier, so the two completions must not share one.
    uint32_t mbar_u32[2];
    uint32_t parity[2];       // phase the next wait on each expects
    unsigned char* a_ptr;     // generic pointer to the A slots
    const unsigned char* tab; // the context's table image in global memory (layout 1)
    int row;                  // this thread's row in the tile, 0..127
    int bar_id;               // named barrier of the tile
    bool leader;              // the thread that issues the tile's MMAs
};

__device__ __forceinline__ uint64_t tc5_desc(uint32_t saddr, uint32_t lbo) {
    // K-major, no swizzle, version 1; 8-row groups 128 bytes apart; lbo = bytes between the two 16-byte K chunks
    return (uint64_t)((saddr & 0x3FFFF) >> 4) | ((uint64_t)(lbo >> 4) << 16) | ((uint64_t)(128 >> 4) << 32) | (1ull << 46);
}
constexpr uint32_t kTc5Idesc = (2u << 4) | ((32u >> 3) << 17) | ((128u >> 4) << 24);   // u8 x u8 -> s32, M 128, N 32

// D[tile columns d_col .. +31] (+)= A(slot) x Tab(b_u32)^T
__device__ __forceinline__ void tc5_mma(const Tc5Tile& t, uint32_t a_slot, uint32_t b_u32, uint32_t d_col, bool accumulate) {
    const uint64_t da = tc5_desc(t.a_u32 + a_slot * kTc5ASlot, 2048), db = tc5_desc(b_u32, 512);
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                 "tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, {%5, %5, %5, %5}, p;\n\t}"
                 ::"r"(t.tmem + d_col), "l"(da), "l"(db), "r"(kTc5Idesc), "r"((uint32_t)accumulate), "r"(0u) : "memory");
}
__device__ __forceinline__ void tc5_commit(const Tc5Tile& t, int which) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(t.mbar_u32[which]) : "memory");
}

// this thread's 32 bytes of a value -> A slot
__device__ __forceinline__ void tc5_store_a(const Tc5Tile& t, uint32_t slot, const uint32_t* x) {
    unsigned char* p = t.a_ptr + slot * kTc5ASlot + t.row * 16;
    *reinterpret_cast<uint4*>(p) = make_uint4(x[0], x[1], x[2], x[3]);
    *reinterpret_cast<uint4*>(p + 2048) = make_uint4(x[4], x[5], x[6], x[7]);
}

// All 128 threads of the tile: what they wrote to shared memory (A slots, staged tables) becomes visible to the tensor
// core, and everyone's TMEM reads of the previous step are complete, before the leader issues.
__device__ __forceinline__ void tc5_step_begin(const Tc5Tile& t) {
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    asm volatile("bar.sync %0, 128;" ::"r"(t.bar_id) : "memory");
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
}

// waits for the next commit of the tile's leader; a wait that does not end is a bug, not a reason to hang the GPU
__device__ __forceinline__ void tc5_wait(Tc5Tile& t, int which) {
    uint32_t ok = 0;
    for (uint32_t spin = 0; spin < (1u << 26) && !ok; spin++) {
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                     : "=r"(ok) : "r"(t.mbar_u32[which]), "r"(t.parity[which]) : "memory");
    }
    if (!ok) __trap();
    t.parity[which] ^= 1;
    __syncwarp();   // the TMEM loads that follow are warp-collective
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
}

// the 32 column sums of this thread's row, tile columns d_col .. +31
__device__ __forceinline__ void tc5_load(const Tc5Tile& t, uint32_t d_col, int (&v)[32]) {
    const uint32_t taddr = t.tmem + d_col + ((uint32_t)(t.row & 96) << 16);   // lane quarter of this warp
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,"
                 "%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
                 : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]),
                   "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]), "=r"(v[16]),
                   "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), "=r"(v[24]),
                   "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
                 : "r"(taddr) : "memory");
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}

// 32 column sums (+ the additive constant K 2^32 mod p, eight limbs at kptr) -> 8 limbs, value < p + 2^241:
// columns 4l, 4l+1 and 4l+2, 4l+3 pair up (sum < 2^32 / 257), the pairs form two 256-bit numbers X + (Y << 16), then one
// 32-bit Montgomery row -- the same sweep as variant 6, minus its exchange through shared memory
template <class F>
__device__ __forceinline__ void tc5_sweep(uint32_t* out, const int (&c)[32], const unsigned char* kptr) {
    uint32_t X[8], Y[8];
#pragma unroll
    for (int l = 0; l < 8; l++) {
        X[l] = (uint32_t)c[4 * l] + ((uint32_t)c[4 * l + 1] << 8);
        Y[l] = (uint32_t)c[4 * l + 2] + ((uint32_t)c[4 * l + 3] << 8);
    }
    uint32_t ys[9];
    ys[0] = Y[0] << 16;
#pragma unroll
    for (int k = 1; k < 8; k++) ys[k] = __funnelshift_l(Y[k - 1], Y[k], 16);
    ys[8] = Y[7] >> 16;
    const uint4 k0 = __ldg(reinterpret_cast<const uint4*>(kptr)), k1 = __ldg(reinterpret_cast<const uint4*>(kptr) + 1);
    uint32_t v[9];
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
        : "r"(X[0]), "r"(X[1]), "r"(X[2]), "r"(X[3]), "r"(X[4]), "r"(X[5]), "r"(X[6]), "r"(X[7]), "r"(ys[8]), "r"(ys[0]),
          "r"(ys[1]), "r"(ys[2]), "r"(ys[3]), "r"(ys[4]), "r"(ys[5]), "r"(ys[6]), "r"(ys[7]));
    asm("add.cc.u32 %0, %0, %9;\n\t"
        "addc.cc.u32 %1, %1, %10;\n\t"
        "addc.cc.u32 %2, %2, %11;\n\t"
        "addc.cc.u32 %3, %3, %12;\n\t"
        "addc.cc.u32 %4, %4, %13;\n\t"
        "addc.cc.u32 %5, %5, %14;\n\t"
        "addc.cc.u32 %6, %6, %15;\n\t"
        "addc.cc.u32 %7, %7, %16;\n\t"
        "addc.u32 %8, %8, 0;"
        : "+r"(v[0]), "+r"(v[1]), "+r"(v[2]), "+r"(v[3]), "+r"(v[4]), "+r"(v[5]), "+r"(v[6]), "+r"(v[7]), "+r"(v[8])
        : "r"(k0.x), "r"(k0.y), "r"(k0.z), "r"(k0.w), "r"(k1.x), "r"(k1.y), "r"(k1.z), "r"(k1.w));
    const uint32_t m = v[0] * np0<F>();
    uint32_t c8 = 0;
    mad4(v, plimb<F, 0>(), plimb<F, 2>(), plimb<F, 4>(), plimb<F, 6>(), m, c8);
    uint32_t x[8] = {v[1], v[2], v[3], v[4], v[5], v[6], v[7], v[8] + c8};
    mad4_nc(x, plimb<F, 1>(), plimb<F, 3>(), plimb<F, 5>(), plimb<F, 7>(), m);
#pragma unroll
    for (int k = 0; k < 8; k++) out[k] = x[k];
}

// n_kb KiB of tables, global -> the tile's staging buffer `buf` at KiB position `at`, by cp.async (no registers held)
__device__ __forceinline__ void tc5_stage(const Tc5Tile& t, uint32_t buf_u32, uint32_t at, const unsigned char* src, uint32_t n_kb) {
    for (uint32_t q = t.row; q < n_kb * 64; q += kTc5TileRows) {
        asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(buf_u32 + at * kTc5Tab + q * 16), "l"(src + q * 16) : "memory");
    }
}
__device__ __forceinline__ void tc5_stage_wait() {
    asm volatile("cp.async.commit_group;" ::: "memory");
    asm volatile("cp.async.wait_group 0;" ::: "memory");
}

// One full round: T S-boxes into A slots 0..T-1, T x T instructions, T sweeps.  mat_u32: the dense matrix's tables in shared
// memory ([row i][column j]), kp: the T additive constants of the round.
template <class F, int T, int ALPHA>
__device__ __forceinline__ void tc5_full_round(uint32_t (&st)[T][8], uint32_t mat_u32, const unsigned char* kp, Tc5Tile& t) {
#pragma unroll
    for (int j = 0; j < T; j++) {
        tc_sbox<F, ALPHA>(st[j]);
        tc5_store_a(t, j, st[j]);
    }
    tc5_step_begin(t);
    if (t.leader) {
#pragma unroll
        for (int i = 0; i < T; i++)
#pragma unroll
            for (int j = 0; j < T; j++) tc5_mma(t, j, mat_u32 + (i * T + j) * kTc5Tab, 32 * i, j > 0);
        tc5_commit(t, 0);
    }
    tc5_wait(t, 0);
#pragma unroll
    for (int i = 0; i < T; i++) {
        int c[32];
        tc5_load(t, 32 * i, c);
        tc5_sweep<F>(st[i], c, kp + i * kTc5KBytes);
    }
}

template <class F, int T, int ALPHA>
__device__ __forceinline__ void permute_tc5(uint32_t (&st)[T][8], const PermDesc& d, const TcDesc& td, Tc5Tile& t) {
    static_assert(ALPHA == 5 || ALPHA == 17, "variant 7: alpha 5 or 17");
    static_assert(T <= 4, "variant 7: a tile's 128 TMEM columns hold T <= 4 output values");
    constexpr uint32_t kRoundK = T * kTc5KBytes, kRoundPart = 2 * (T - 1) * kTc5Tab, kPairTabs = tc5_pair_tables(T);
    // A slots: 0 = y0, 1 = y1, 2.. = lanes 1..T-1 (full rounds use 0..T-1)
#pragma unroll
    for (int j = 0; j < T; j++) {
        uint32_t c[8];
        cload8(c, d.ark_off + j * 8);
        fp_add<F>(st[j], st[j], c);
    }
    const uint32_t half = d.half_full, rp = d.partial;
    const unsigned char* kp = t.tab + td.k_off;
    const uint32_t mat_u32 = t.static_u32, pre_u32 = t.static_u32 + T * T * kTc5Tab, ident_u32 = t.static_u32 + 2 * T * T * kTc5Tab;

#pragma unroll 1
    for (uint32_t r = 0; r < half; r++) {
        tc5_full_round<F, T, ALPHA>(st, (r + 1 == half && rp) ? pre_u32 : mat_u32, kp, t);
        kp += kRoundK;
    }

    if (rp) {
        const unsigned char* part = t.tab + td.part_off;
        const unsigned char* pair = t.tab + td.pair_off;
        const uint32_t pairs = rp >> 1;
        uint32_t buf = 0;   // staging buffer in use
        // lanes into their slots; the first round's (or pair's) tables
#pragma unroll
        for (int j = 1; j < T; j++) tc5_store_a(t, 1 + j, st[j]);
        if (rp & 1) {
            tc5_stage(t, t.tab_u32, 0, part, 2 * (T - 1));
        } else {
            tc5_stage(t, t.tab_u32, 0, part, 4 * (T - 1));
            tc5_stage(t, t.tab_u32, 4 * (T - 1), pair, 1);
        }
        tc_sbox<F, ALPHA>(st[0]);
        tc5_store_a(t, 0, st[0]);
        tc5_stage_wait();
        if (rp & 1) {
            // an odd number of partial rounds: the first one on its own.  tables: w_1.. at 0, v_1.. at T-1
            const uint32_t tb = t.tab_u32;
            if (pairs) {   // the first pair's tables go to the other buffer meanwhile
                tc5_stage(t, t.tab_u32 + kPairTabs * kTc5Tab, 0, part + kRoundPart, 4 * (T - 1));
                tc5_stage(t, t.tab_u32 + kPairTabs * kTc5Tab, 4 * (T - 1), pair, 1);
            }
            tc5_step_begin(t);
            if (t.leader) {
                tc5_mma(t, 0, mat_u32, 0, false);
#pragma unroll
                for (int j = 1; j < T; j++) tc5_mma(t, 1 + j, tb + (j - 1) * kTc5Tab, 0, true);
                tc5_commit(t, 0);
#pragma unroll
                for (int j = 1; j < T; j++) {
                    tc5_mma(t, 0, tb + (T - 1 + j - 1) * kTc5Tab, 32 * j, false);
                    tc5_mma(t, 1 + j, ident_u32, 32 * j, true);
                }
                tc5_commit(t, 1);
            }
            tc5_wait(t, 0);
            {
                int c[32];
                tc5_load(t, 0, c);
                tc5_sweep<F>(st[0], c, kp);
            }
            if (pairs) tc_sbox<F, ALPHA>(st[0]);
            tc5_wait(t, 1);
#pragma unroll
            for (int j = 1; j < T; j++) {
                int c[32];
                tc5_load(t, 32 * j, c);
                tc5_sweep<F>(st[j], c, kp + j * kTc5KBytes);
                if (pairs) tc5_store_a(t, 1 + j, st[j]);
            }
            if (pairs) {
                tc5_store_a(t, 0, st[0]);
                tc5_stage_wait();
            }
            kp += kRoundK;
            part += kRoundPart;
            buf = 1;
        }
#pragma unroll 1
        for (uint32_t q = 0; q < pairs; q++) {
            const bool last = q + 1 == pairs;
            // tables of this pair: w(i) at 0, v(i) at T-1, w(i+1) at 2(T-1), v(i+1) at 3(T-1), c_q at 4(T-1)
            const uint32_t tb = t.tab_u32 + buf * kPairTabs * kTc5Tab;
            // first row 0:  M00 y0 + sum_j w_j(i) s_j
            tc5_step_begin(t);
            if (t.leader) {
                tc5_mma(t, 0, mat_u32, 0, false);
#pragma unroll
                for (int j = 1; j < T; j++) tc5_mma(t, 1 + j, tb + (j - 1) * kTc5Tab, 0, true);
                tc5_commit(t, 0);
            }
            // the next pair's tables start their way into the other buffer (its last readers finished a pair ago)
            if (!last) {
                tc5_stage(t, t.tab_u32 + (buf ^ 1) * kPairTabs * kTc5Tab, 0, part + 2 * kRoundPart, 4 * (T - 1));
                tc5_stage(t, t.tab_u32 + (buf ^ 1) * kPairTabs * kTc5Tab, 4 * (T - 1), pair + kTc5Tab, 1);
            }
            tc5_wait(t, 0);
            {
                int c[32];
                tc5_load(t, 0, c);
                tc5_sweep<F>(st[0], c, kp);
            }
            tc_sbox<F, ALPHA>(st[0]);
            tc5_store_a(t, 1, st[0]);
            // second row 0:  M00 y1 + c_q y0 + sum_j w_j(i+1) s_j ; then the lanes:  1 s_j + v_j(i) y0 + v_j(i+1) y1
            tc5_step_begin(t);
            if (t.leader) {
                tc5_mma(t, 1, mat_u32, 0, false);
                tc5_mma(t, 0, tb + 4 * (T - 1) * kTc5Tab, 0, true);
#pragma unroll
                for (int j = 1; j < T; j++) tc5_mma(t, 1 + j, tb + (2 * (T - 1) + j - 1) * kTc5Tab, 0, true);
                tc5_commit(t, 0);
#pragma unroll
                for (int j = 1; j < T; j++) {
                    tc5_mma(t, 1, tb + (3 * (T - 1) + j - 1) * kTc5Tab, 32 * j, false);
                    tc5_mma(t, 0, tb + (T - 1 + j - 1) * kTc5Tab, 32 * j, true);
                    tc5_mma(t, 1 + j, ident_u32, 32 * j, true);
                }
                tc5_commit(t, 1);
            }
            tc5_wait(t, 0);
            {
                int c[32];
                tc5_load(t, 0, c);
                tc5_sweep<F>(st[0], c, kp + kRoundK);
            }
            // independent of each other: the S-box of the next pair's first round and this pair's lane sweeps
            if (!last) tc_sbox<F, ALPHA>(st[0]);
            tc5_wait(t, 1);
#pragma unroll
            for (int j = 1; j < T; j++) {
                int c[32];
                tc5_load(t, 32 * j, c);
                tc5_sweep<F>(st[j], c, kp + kRoundK + j * kTc5KBytes);
                if (!last) tc5_store_a(t, 1 + j, st[j]);
            }
            if (!last) {
                tc5_store_a(t, 0, st[0]);
                tc5_stage_wait();
            }
            kp += 2 * kRoundK;
            part += 2 * kRoundPart;
            pair += kTc5Tab;
            buf ^= 1;
        }
    }

#pragma unroll 1
    for (uint32_t r = half + rp; r < d.total; r++) {
        tc5_full_round<F, T, ALPHA>(st, mat_u32, kp, t);
        kp += kRoundK;
    }

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
