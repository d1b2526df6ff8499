// VARIANT 7: variant 6's contraction (products by round constants as u8 x u8 -> s32 matrix products against byte-digit
// tables, see poseidon_tc.cuh) on the 5th-generation tensor cores: tcgen05.mma.kind::i8 with both operands in shared
// memory and the accumulators in TMEM -- and, because a Montgomery reduction is itself a product by constants
// (z mod p = sum_k byte_k(z) (256^k mod p)), every reduction of the S-box as well.
//
// A thread block is three to five TILES of 128 sponges (4 warps each; 5 up to t = 3, 4 at t = 4, 3 at t = 5).  Per tile
//   * every thread writes the bytes of a value ONCE into an A slot of the tile (K-major core-matrix layout: 16-byte chunk c
//     of row r at c * 2048 + r * 16): 32 bytes for a state element, 64 bytes for an UNREDUCED product of two elements;
//   * ONE thread issues the MMAs of a step (M = 128 sponges, K = 32 bytes per instruction, N = 32 columns per constant; the
//     constants that multiply the same value are stacked along N, so one instruction pair serves row 0 AND all lanes) and
//     commits them to the tile's mbarrier; measured on B200: 67 cycles per instruction per SM for every N <= 128, so the
//     instruction COUNT is what the layout minimises;
//   * the accumulators live in the tile's 32 t TMEM columns, and tcgen05.ld.32x32b hands thread r the 32 column sums of ITS
//     OWN row -- no exchange, no fragments;
//   * the sweep (pair-combine, carry add, one 32-bit Montgomery row) turns 32 column sums into a value < p + 2^249.
// The S-box x^5 is then: square (36 wide multiplies, unreduced) -> tensor-core reduction -> square -> reduction -> product
// with x (64 wide multiplies, unreduced) -> straight into the linear layer, whose tables carry the extra R^-1.  No
// Montgomery reduction runs on the integer pipe: 136 wide multiplies per S-box instead of 328, for three more trips
// through the tensor core per round, which the other warps of the SM hide.
//
// Shared memory (behind the staged constant image): [M00 table, ident, reduction table][A slots per tile][U], where U holds
// the dense matrices (M, M_pre, stacked by column) during the full rounds and the per-tile table staging buffers during the
// partial rounds; the matrices are re-read from the L2-resident image after the partial rounds.
#pragma once
#include "poseidon_tc.cuh"

namespace psd {

constexpr uint32_t kTc5Tab = 1024;      // one 32-row table: [k / 16][column 0..31][k % 16]
constexpr uint32_t kTc5Chunk = 2048;    // one 16-byte K chunk of an A operand: [row 0..127][16 bytes]
constexpr uint32_t kTc5Narrow = 4096;   // A slot of a 32-byte value
constexpr uint32_t kTc5Wide = 8192;     // A slot of a 64-byte value
constexpr uint32_t kTc5TileRows = 128;

#if defined(__CUDACC__)
#define PSD_HD __host__ __device__
#else
#define PSD_HD
#endif
// block-wide small tables: M00 (64 rows) at 0, ident (32 rows) at 2 KiB, the reduction table (64 rows) at 3 KiB
PSD_HD constexpr uint32_t tc5_small_bytes() { return 5 * 1024; }
// A slots of a tile: full rounds T wide slots; partial rounds y0, y1 wide and T - 1 narrow lanes
PSD_HD constexpr uint32_t tc5_a_bytes(uint32_t t) {
    return t * kTc5Wide > 2 * kTc5Wide + (t - 1) * kTc5Narrow ? t * kTc5Wide : 2 * kTc5Wide + (t - 1) * kTc5Narrow;
}
// staging of a tile for one pair of partial rounds: w(i) [t-1], the stack (c_q, v(i)) [2t], w(i+1) [t-1], the stack
// (M00, v(i+1)) [2t] KiB
PSD_HD constexpr uint32_t tc5_stage_bytes(uint32_t t) { return (6 * t - 2) * 1024; }
// U: both dense matrices (M and M_pre, 2 t^2 KiB each) up to t = 4, one at a time at t = 5 -- or the tiles' staging buffers
PSD_HD constexpr bool tc5_both_matrices(uint32_t t) { return t <= 4; }
PSD_HD constexpr uint32_t tc5_u_bytes(uint32_t t, uint32_t tiles) {
    return (tc5_both_matrices(t) ? 4 : 2) * t * t * 1024 > tiles * tc5_stage_bytes(t) ? (tc5_both_matrices(t) ? 4 : 2) * t * t * 1024
                                                                                       : tiles * tc5_stage_bytes(t);
}
// TMEM columns a block of `tiles` tiles allocates: 32 t per tile, rounded up to a power of two >= 32
PSD_HD constexpr uint32_t tc5_tmem_cols(uint32_t t, uint32_t tiles) {
    uint32_t c = 32;
    while (c < 32 * t * tiles) c *= 2;
    return c;
}
PSD_HD constexpr uint32_t tc5_smem_bytes(uint32_t t, uint32_t tiles) {
    return tc5_small_bytes() + tiles * tc5_a_bytes(t) + tc5_u_bytes(t, tiles) + 128;
}

#if defined(__CUDACC__)

struct Tc5Tile {
    uint32_t tmem;            // TMEM address of this tile's column 0 (lane field 0)
    uint32_t a_q;             // shared-memory address / 16 of the tile's A slots
    uint32_t stage_u32;       // shared-memory address of the tile's staging buffer (inside U)
    uint32_t small_q;         // address / 16 of the block's small tables
    uint32_t u_u32;           // address of U
    // Two mbarriers per tile: a step commits twice (row 0 first, the lanes behind it), and a waiter may be at most one
    // phase behind an mbarrier, so the two completions must not share one.
    uint32_t mbar_u32[2];
    uint32_t parity[2];       // phase the next wait on each expects
    unsigned char* a_ptr;     // generic pointer to the A slots
    const unsigned char* tab; // the context's table image in global memory (layout 1)
    int row;                  // this thread's row in the tile, 0..127
    int bar_id;               // named barrier of the tile
    bool leader_warp;         // warp-uniform: the tile's first warp, one elected thread of which issues the tile's MMAs
};

// Shared-memory operand addresses travel as 16-byte units (`q` = address / 16: every operand here is 16-byte aligned and the
// shared window is below 256 KiB), so a descriptor is one 32-bit add away from its region's base: the issuing warp carries
// 15 % more instructions than its three siblings, and they wait for it at every step.
__device__ __forceinline__ uint64_t tc5_desc(uint32_t q, uint32_t lbo) {
    // K-major, no swizzle, version 1; 8-row groups 128 bytes apart; lbo = bytes between the two 16-byte K chunks
    const uint32_t lo = q + ((lbo >> 4) << 16), hi = (128u >> 4) | (1u << 14);
    uint64_t d;
    asm("mov.b64 %0, {%1, %2};" : "=l"(d) : "r"(lo), "r"(hi));
    return d;
}
// u8 x u8 -> s32, M 128, N n
__device__ __forceinline__ constexpr uint32_t tc5_idesc(uint32_t n) { return (2u << 4) | ((n >> 3) << 17) | ((128u >> 4) << 24); }

// D[columns d .. d + 32 blocks - 1] (+)= A (128 rows x 32 bytes at 16 a_q) x B^T (32 blocks columns x 32 bytes at 16 b_q)
__device__ __forceinline__ void tc5_mma(uint32_t d, uint32_t a_q, uint32_t b_q, uint32_t blocks, bool accumulate) {
    const uint64_t da = tc5_desc(a_q, kTc5Chunk), db = tc5_desc(b_q, blocks * 512);
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                 "tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, {%5, %5, %5, %5}, p;\n\t}"
                 ::"r"(d), "l"(da), "l"(db), "r"(tc5_idesc(32 * blocks)), "r"((uint32_t)accumulate), "r"(0u) : "memory");
}
// a 64-byte value against a stack of `blocks` 64-row tables: two instructions
__device__ __forceinline__ void tc5_mma_wide(uint32_t d, uint32_t a_q, uint32_t b_q, uint32_t blocks, bool accumulate) {
    tc5_mma(d, a_q, b_q, blocks, accumulate);
    tc5_mma(d, a_q + 2 * kTc5Chunk / 16, b_q + 2 * blocks * 512 / 16, blocks, true);
}
__device__ __forceinline__ void tc5_commit(const Tc5Tile& t, int which) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(t.mbar_u32[which]) : "memory");
}

// One thread of a converged warp.  The MMAs are issued from warp-uniform control flow with warp-uniform operands (the tile
// index comes through a shuffle, see sponge_kernel.cuh): behind a per-thread predicate ptxas wraps every instruction in a
// broadcast loop that ALL warps execute (measured: 11 % of the kernel's instructions).
__device__ __forceinline__ bool tc5_elect() {
    uint32_t pred = 0;
    asm volatile("{\n\t.reg .pred p;\n\telect.sync _|p, 0xffffffff;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(pred));
    return pred != 0;
}

// this thread's bytes of a value -> A slot at byte offset `off` of the tile's A region
__device__ __forceinline__ void tc5_store_a(const Tc5Tile& t, uint32_t off, const uint32_t* x) {
    unsigned char* p = t.a_ptr + off + t.row * 16;
    *reinterpret_cast<uint4*>(p) = make_uint4(x[0], x[1], x[2], x[3]);
    *reinterpret_cast<uint4*>(p + kTc5Chunk) = make_uint4(x[4], x[5], x[6], x[7]);
}
__device__ __forceinline__ void tc5_store_wide(const Tc5Tile& t, uint32_t off, const uint32_t* x) {
    tc5_store_a(t, off, x);
    tc5_store_a(t, off + 2 * kTc5Chunk, x + 8);
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
        // (measured and left out: a suspend-time hint here -- 1 poll per wait instead of 3.6, 1.5 % slower; an arrive-only
        // mbarrier in place of the tile barrier of tc5_step_begin -- 1 % slower)
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

// 32 column sums (+ the additive constant K 2^32 mod p, eight limbs at kptr, if HASK) -> 8 limbs, value < p + 2^249:
// columns 4l, 4l+1 and 4l+2, 4l+3 pair up (sum < 2^32 / 257), the pairs form two 256-bit numbers X + (Y << 16), then one
// 32-bit Montgomery row -- the same sweep as variant 6, minus its exchange through shared memory
// BIG: column sums up to 2^31 (the full rounds at t = 5 add ten 32-product units): the four columns of a limb are combined
// in 64 bits instead.
template <class F, bool HASK, bool BIG = false>
__device__ __forceinline__ void tc5_sweep(uint32_t* out, const int (&c)[32], const unsigned char* kptr) {
    uint32_t X[8], ys[9];
    if (BIG) {
        ys[0] = 0;
#pragma unroll
        for (int l = 0; l < 8; l++) {
            const uint64_t w = (uint64_t)(uint32_t)c[4 * l] + ((uint64_t)(uint32_t)c[4 * l + 1] << 8) +
                               ((uint64_t)(uint32_t)c[4 * l + 2] << 16) + ((uint64_t)(uint32_t)c[4 * l + 3] << 24);
            X[l] = (uint32_t)w;
            ys[l + 1] = (uint32_t)(w >> 32);
        }
    } else {
        uint32_t Y[8];
#pragma unroll
        for (int l = 0; l < 8; l++) {
            X[l] = (uint32_t)c[4 * l] + ((uint32_t)c[4 * l + 1] << 8);
            Y[l] = (uint32_t)c[4 * l + 2] + ((uint32_t)c[4 * l + 3] << 8);
        }
        ys[0] = Y[0] << 16;
#pragma unroll
        for (int k = 1; k < 8; k++) ys[k] = __funnelshift_l(Y[k - 1], Y[k], 16);
        ys[8] = Y[7] >> 16;
    }
    if (HASK) {
        // the additive constant joins the shifted half: ys += K (ys[8] stays small: it only takes the carry)
        const uint4 k0 = __ldg(reinterpret_cast<const uint4*>(kptr)), k1 = __ldg(reinterpret_cast<const uint4*>(kptr) + 1);
        asm("add.cc.u32 %0, %0, %9;\n\t"
            "addc.cc.u32 %1, %1, %10;\n\t"
            "addc.cc.u32 %2, %2, %11;\n\t"
            "addc.cc.u32 %3, %3, %12;\n\t"
            "addc.cc.u32 %4, %4, %13;\n\t"
            "addc.cc.u32 %5, %5, %14;\n\t"
            "addc.cc.u32 %6, %6, %15;\n\t"
            "addc.cc.u32 %7, %7, %16;\n\t"
            "addc.u32 %8, %8, 0;"
            : "+r"(ys[0]), "+r"(ys[1]), "+r"(ys[2]), "+r"(ys[3]), "+r"(ys[4]), "+r"(ys[5]), "+r"(ys[6]), "+r"(ys[7]), "+r"(ys[8])
            : "r"(k0.x), "r"(k0.y), "r"(k0.z), "r"(k0.w), "r"(k1.x), "r"(k1.y), "r"(k1.z), "r"(k1.w));
    }
    // (X + ys + m p) / 2^32 without ever forming X + ys: the Montgomery row's even products accumulate onto X's register
    // pairs, its odd products onto the pairs (ys1, ys2), (ys3, ys4), .. -- both are pairs as IMAD.WIDE wants them -- and one
    // 9-limb add joins the two halves one limb apart.  (Accumulating both product chains onto ONE sum needs the sum in two
    // pairings, i.e. eight register moves per sweep.)  Limb 0 of the total is zero by the choice of m: its add only makes the
    // carry.
    const uint32_t m = (X[0] + ys[0]) * np0<F>();
    uint32_t ca = 0;
    // (redc_lo / redc_hi: BLS12-381's p0 = 1 and p1 = 2^32 - 1 turn two of the eight products into adds)
    redc_lo<F>(X, m, ca);
    if (F::kBits == 255) {
        uint32_t none = 0;   // the top pair cannot carry out: ys[8] only holds a carry and m p7 < 2^63
        redc_hi<F>(ys + 1, m, none);
    } else {
        mad4_nc(ys + 1, plimb<F, 1>(), plimb<F, 3>(), plimb<F, 5>(), plimb<F, 7>(), m);
    }
    uint32_t zero;
    asm("add.cc.u32 %8, %9, %18;\n\t"
        "addc.cc.u32 %0, %10, %19;\n\t"
        "addc.cc.u32 %1, %11, %20;\n\t"
        "addc.cc.u32 %2, %12, %21;\n\t"
        "addc.cc.u32 %3, %13, %22;\n\t"
        "addc.cc.u32 %4, %14, %23;\n\t"
        "addc.cc.u32 %5, %15, %24;\n\t"
        "addc.cc.u32 %6, %16, %25;\n\t"
        "addc.u32 %7, %17, %26;"
        : "=r"(out[0]), "=r"(out[1]), "=r"(out[2]), "=r"(out[3]), "=r"(out[4]), "=r"(out[5]), "=r"(out[6]), "=r"(out[7]), "=r"(zero)
        : "r"(X[0]), "r"(X[1]), "r"(X[2]), "r"(X[3]), "r"(X[4]), "r"(X[5]), "r"(X[6]), "r"(X[7]), "r"(ca), "r"(ys[0]), "r"(ys[1]),
          "r"(ys[2]), "r"(ys[3]), "r"(ys[4]), "r"(ys[5]), "r"(ys[6]), "r"(ys[7]), "r"(ys[8]));
    (void)zero;
}

// `bytes` of tables, global -> shared at dst_u32, by cp.async (no registers held), spread over the tile's threads
__device__ __forceinline__ void tc5_stage(const Tc5Tile& t, uint32_t dst_u32, const unsigned char* src, uint32_t bytes) {
    for (uint32_t q = t.row; q < bytes / 16; q += kTc5TileRows) {
        asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst_u32 + q * 16), "l"(src + q * 16) : "memory");
    }
}
__device__ __forceinline__ void tc5_stage_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
__device__ __forceinline__ void tc5_stage_wait() { asm volatile("cp.async.wait_group 0;" ::: "memory"); }

// The dense matrices' stacks, image -> U, by the whole block; ends with everything visible to the tensor core.
// first_off: the matrix that goes to the front of U; second_off: the one behind it, or 0 for none.
template <int T>
__device__ __forceinline__ void tc5_load_matrices(const Tc5Tile& t, uint32_t first_off, uint32_t second_off) {
    constexpr uint32_t kMat = 2 * T * T * 1024;
    for (uint32_t q = threadIdx.x; q < kMat / 16; q += blockDim.x) {
        asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(t.u_u32 + q * 16), "l"(t.tab + first_off + q * 16) : "memory");
        if (second_off)
            asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(t.u_u32 + kMat + q * 16), "l"(t.tab + second_off + q * 16) : "memory");
    }
    tc5_stage_commit();
    tc5_stage_wait();
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    __syncthreads();
}

// the S-box of ONE value: x -> the unreduced 64 bytes of x^ALPHA in the wide slot at `off`.
// LANES: the lane columns of the previous step (tile columns 32 j, complete) are swept and written back to their A slots
// first -- in one straight line with the first squaring, which depends on none of it, so that the compiler can fill the
// squaring's multiplier latencies with the sweeps' ALU work (+0.5 % on C2 against doing it while the first reduction is on
// its way through the tensor core).
template <class F, int T, int ALPHA, bool LANES>
__device__ __forceinline__ void tc5_sbox_slot(const uint32_t* x, uint32_t off, Tc5Tile& t, uint32_t lane_off = 0) {
    uint32_t y[8], in[8], w[16];
#pragma unroll
    for (int k = 0; k < 8; k++) in[k] = y[k] = x[k];
    if (LANES) {
#pragma unroll
        for (int j = 1; j < T; j++) {
            int c[32];
            uint32_t lane[8];
            tc5_load(t, 32 * j, c);
            tc5_sweep<F, false>(lane, c, nullptr);
            tc5_store_a(t, lane_off + (j - 1) * kTc5Narrow, lane);
        }
    }
#pragma unroll
    for (int q = 0; q < (ALPHA == 5 ? 2 : 4); q++) {
        sqr_wide(w, y);
        tc5_store_wide(t, off, w);
        tc5_step_begin(t);
        if (t.leader_warp && tc5_elect()) {
            tc5_mma_wide(t.tmem, t.a_q + off / 16, t.small_q + 3 * 1024 / 16, 1, false);
            tc5_commit(t, 0);
        }
        tc5_wait(t, 0);
        int c[32];
        tc5_load(t, 0, c);
        tc5_sweep<F, false>(y, c, nullptr);
    }
    mul_wide(w, y, in);
    tc5_store_wide(t, off, w);
}

// One full round: T S-boxes into the wide slots 0..T-1, 2 T instructions (S-box output j against column j of the matrix,
// all rows stacked), T sweeps.  mat_q: the matrix's stacks in U (address / 16), kp: the T additive constants of the round.
template <class F, int T, int ALPHA>
__device__ __forceinline__ void tc5_full_round(uint32_t (&st)[T][8], uint32_t mat_q, const unsigned char* kp, Tc5Tile& t) {
    uint32_t y[T][8];
#pragma unroll
    for (int j = 0; j < T; j++)
#pragma unroll
        for (int k = 0; k < 8; k++) y[j][k] = st[j][k];
#pragma unroll
    for (int q = 0; q < (ALPHA == 5 ? 2 : 4); q++) {
#pragma unroll
        for (int j = 0; j < T; j++) {
            uint32_t w[16];
            sqr_wide(w, y[j]);
            tc5_store_wide(t, j * kTc5Wide, w);
        }
        tc5_step_begin(t);
        if (t.leader_warp && tc5_elect()) {
#pragma unroll
            for (int j = 0; j < T; j++) tc5_mma_wide(t.tmem + 32 * j, t.a_q + j * kTc5Wide / 16, t.small_q + 3 * 1024 / 16, 1, false);
            tc5_commit(t, 0);
        }
        tc5_wait(t, 0);
#pragma unroll
        for (int j = 0; j < T; j++) {
            int c[32];
            tc5_load(t, 32 * j, c);
            tc5_sweep<F, false>(y[j], c, nullptr);
        }
    }
#pragma unroll
    for (int j = 0; j < T; j++) {
        uint32_t w[16];
        mul_wide(w, y[j], st[j]);
        tc5_store_wide(t, j * kTc5Wide, w);
    }
    tc5_step_begin(t);
    if (t.leader_warp && tc5_elect()) {
#pragma unroll
        for (int j = 0; j < T; j++) tc5_mma_wide(t.tmem, t.a_q + j * kTc5Wide / 16, mat_q + j * 2 * T * 1024 / 16, T, j > 0);
        tc5_commit(t, 0);
    }
    tc5_wait(t, 0);
#pragma unroll
    for (int i = 0; i < T; i++) {
        int c[32];
        tc5_load(t, 32 * i, c);
        tc5_sweep<F, true, (T > 4)>(st[i], c, kp + i * kTc5KBytes);
    }
}

template <class F, int T, int ALPHA>
__device__ __forceinline__ void permute_tc5(uint32_t (&st)[T][8], const PermDesc& d, const TcDesc& td, Tc5Tile& t) {
    static_assert(ALPHA == 5 || ALPHA == 17, "variant 7: alpha 5 or 17");
    static_assert(T >= 2 && T <= 5, "variant 7: widths 2..5 (a tile owns 32 T of the SM's 512 TMEM columns)");
    constexpr bool kBoth = tc5_both_matrices(T);
    constexpr uint32_t N = T - 1, kRoundK = T * kTc5KBytes, kStack = 2 * T * 1024, kRoundPart = N * kTc5Tab + kStack;
    // A slots of the partial rounds
    constexpr uint32_t kY0 = 0, kY1 = kTc5Wide, kLane = 2 * kTc5Wide;
    // staging: w(i), the stack (c_q, v(i)), then -- one contiguous piece of the image -- w(i+1) and the stack (M00, v(i+1))
    const uint32_t sWa = t.stage_u32, sBy0 = sWa + N * kTc5Tab, sWb = sBy0 + kStack;
    // the same as MMA operands (address / 16), and the A slots
    const uint32_t qWa = sWa >> 4, qBy0 = sBy0 >> 4, qWb = sWb >> 4, qBy1 = qWb + N * kTc5Tab / 16;
    const uint32_t qM00 = t.small_q, qIdent = t.small_q + 2 * 1024 / 16;
    const uint32_t qY0 = t.a_q + kY0 / 16, qY1 = t.a_q + kY1 / 16, qLane = t.a_q + kLane / 16;
#pragma unroll
    for (int j = 0; j < T; j++) {
        uint32_t c[8];
        cload8(c, d.ark_off + j * 8);
        fp_add<F>(st[j], st[j], c);
    }
    const uint32_t half = d.half_full, rp = d.partial;
    const unsigned char* kp = t.tab + td.k_off;
    // on entry U holds M (and M_pre behind it where both fit)
    const uint32_t mat_q = t.u_u32 >> 4, pre_q = kBoth ? mat_q + T * kStack / 16 : mat_q;

#pragma unroll 1
    for (uint32_t r = 0; r < half; r++) {
        const bool pre = r + 1 == half && rp;
        if (!kBoth && pre) {
            __syncthreads();   // every tile is through with M
            tc5_load_matrices<T>(t, td.pre_off, 0);
        }
        tc5_full_round<F, T, ALPHA>(st, pre ? pre_q : mat_q, kp, t);
        kp += kRoundK;
    }

    if (rp) {
        const unsigned char* part = t.tab + td.part_off;
        const unsigned char* pair = t.tab + td.pair_off;
        const uint32_t pairs = rp >> 1;
        // every tile has seen its last full round complete: U turns into the staging buffers
        __syncthreads();
        if (rp & 1) {
            tc5_stage(t, sWb, part, kRoundPart);
        } else {
            tc5_stage(t, sWa, part, N * kTc5Tab);
            tc5_stage(t, sBy0, pair, kStack);
            tc5_stage(t, sWb, part + kRoundPart, kRoundPart);
        }
        tc5_stage_commit();
#pragma unroll
        for (int j = 1; j < T; j++) tc5_store_a(t, kLane + (j - 1) * kTc5Narrow, st[j]);
        if (!(rp & 1)) {
            // every pair finds the lanes as column sums in TMEM (the previous step's), the first one too: 1 s_j through the
            // tensor core
            tc5_step_begin(t);
            if (t.leader_warp && tc5_elect()) {
#pragma unroll
                for (int j = 1; j < T; j++) tc5_mma(t.tmem + 32 * j, t.a_q + (kLane + (j - 1) * kTc5Narrow) / 16, t.small_q + 2 * 1024 / 16, 1, false);
                tc5_commit(t, 1);
            }
            tc5_wait(t, 1);
        }
        if (rp & 1) {
            // an odd number of partial rounds: the first one on its own.  y0 against the stack (M00, v): row 0 and lanes
            tc5_sbox_slot<F, T, ALPHA, false>(st[0], kY0, t);
            tc5_stage_wait();
            tc5_step_begin(t);
            if (t.leader_warp && tc5_elect()) {
                tc5_mma_wide(t.tmem, qY0, qBy1, T, false);
#pragma unroll
                for (int j = 1; j < T; j++) tc5_mma(t.tmem, qLane + (j - 1) * kTc5Narrow / 16, qWb + (j - 1) * kTc5Tab / 16, 1, true);
                tc5_commit(t, 0);
#pragma unroll
                for (int j = 1; j < T; j++) tc5_mma(t.tmem + 32 * j, qLane + (j - 1) * kTc5Narrow / 16, qIdent, 1, true);
                tc5_commit(t, 1);
            }
            tc5_wait(t, 0);
            {
                int c[32];
                tc5_load(t, 0, c);
                tc5_sweep<F, true>(st[0], c, kp);
            }
            tc5_wait(t, 1);
            if (pairs) {
                tc5_stage(t, sWa, part + kRoundPart, N * kTc5Tab);
                tc5_stage(t, sBy0, pair, kStack);
                tc5_stage(t, sWb, part + 2 * kRoundPart, kRoundPart);
                tc5_stage_commit();
            } else {
#pragma unroll
                for (int j = 1; j < T; j++) {
                    int c[32];
                    tc5_load(t, 32 * j, c);
                    tc5_sweep<F, true>(st[j], c, kp + j * kTc5KBytes);
                }
            }
            kp += kRoundK;
            part += kRoundPart;
        }
#pragma unroll 1
        for (uint32_t q = 0; q < pairs; q++) {
            const bool last = q + 1 == pairs;
            // first round of the pair:  row 0 = M00 y0 + sum_j w_j(i) s_j
            tc5_sbox_slot<F, T, ALPHA, true>(st[0], kY0, t, kLane);
            tc5_stage_wait();
            tc5_step_begin(t);
            if (t.leader_warp && tc5_elect()) {
                tc5_mma_wide(t.tmem, qY0, qM00, 1, false);
#pragma unroll
                for (int j = 1; j < T; j++) tc5_mma(t.tmem, qLane + (j - 1) * kTc5Narrow / 16, qWa + (j - 1) * kTc5Tab / 16, 1, true);
                tc5_commit(t, 0);
            }
            tc5_wait(t, 0);
            {
                int c[32];
                tc5_load(t, 0, c);
                tc5_sweep<F, true>(st[0], c, kp);
            }
            // second round:  y1 (M00, v(i+1)) + y0 (c_q, v(i)) fill row 0 and the lanes' columns at once; the lanes add
            // w_j(i+1) s_j to row 0 and 1 s_j to their own columns
            tc5_sbox_slot<F, T, ALPHA, false>(st[0], kY1, t);
            tc5_step_begin(t);
            if (t.leader_warp && tc5_elect()) {
                tc5_mma_wide(t.tmem, qY1, qBy1, T, false);
                tc5_mma_wide(t.tmem, qY0, qBy0, T, true);
#pragma unroll
                for (int j = 1; j < T; j++) tc5_mma(t.tmem, qLane + (j - 1) * kTc5Narrow / 16, qWb + (j - 1) * kTc5Tab / 16, 1, true);
                tc5_commit(t, 0);
#pragma unroll
                for (int j = 1; j < T; j++) tc5_mma(t.tmem + 32 * j, qLane + (j - 1) * kTc5Narrow / 16, qIdent, 1, true);
                tc5_commit(t, 1);
            }
            tc5_wait(t, 0);
            {
                int c[32];
                tc5_load(t, 0, c);
                tc5_sweep<F, true>(st[0], c, kp + kRoundK);
            }
            // every instruction that read the A slots and the staging buffers has completed
            tc5_wait(t, 1);
            if (last) {
                // the last partial round's lanes receive the constants moved behind the partial rounds
#pragma unroll
                for (int j = 1; j < T; j++) {
                    int c[32];
                    tc5_load(t, 32 * j, c);
                    tc5_sweep<F, true>(st[j], c, kp + kRoundK + j * kTc5KBytes);
                }
            } else {
                // the next pair's tables; its first S-box sweeps this pair's lanes
                tc5_stage(t, sWa, part + 2 * kRoundPart, N * kTc5Tab);
                tc5_stage(t, sBy0, pair + kStack, kStack);
                tc5_stage(t, sWb, part + 3 * kRoundPart, kRoundPart);
                tc5_stage_commit();
            }
            kp += 2 * kRoundK;
            part += 2 * kRoundPart;
            pair += kStack;
        }
        // all tiles are through with the staging buffers: the matrices return to U
        __syncthreads();
        tc5_load_matrices<T>(t, td.mat_off, kBoth ? td.pre_off : 0);
    }

#pragma unroll 1
    for (uint32_t r = half + rp; r < d.total; r++) {
        tc5_full_round<F, T, ALPHA>(st, mat_q, kp, t);
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
