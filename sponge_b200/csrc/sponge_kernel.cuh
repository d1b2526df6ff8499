// The sponge kernel template (sm_100a): one thread = one sponge; the t-element state lives in
// registers for the whole call sequence.  Instantiated per (field, t) by tu_sponge.cu.
#pragma once
#include <cuda_runtime.h>

#include "kernels.cuh"
#include "poseidon_fp52.cuh"
#include "poseidon_tc5.cuh"

namespace psd {

// ------------------------------------------------------------------------------------------------
// Element load/store.  AoS: elem(i,k) at u32 offset (i*w + k)*8.  SoA: half h of elem(i,k) at u32
// offset ((k*2+h)*n + i)*4, so a warp touches 32 consecutive 16-byte words.
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ void load_elem(uint32_t* x, const uint32_t* base, int layout, uint64_t n,
                                          uint64_t i, uint32_t w, uint32_t k) {
    const uint4* p0;
    const uint4* p1;
    if (layout == kLayoutAos) {
        p0 = reinterpret_cast<const uint4*>(base + (i * w + k) * 8);
        p1 = p0 + 1;
    } else {
        p0 = reinterpret_cast<const uint4*>(base + ((uint64_t)(k * 2) * n + i) * 4);
        p1 = reinterpret_cast<const uint4*>(base + ((uint64_t)(k * 2 + 1) * n + i) * 4);
    }
    const uint4 a = __ldg(p0);
    const uint4 b = __ldg(p1);
    x[0] = a.x; x[1] = a.y; x[2] = a.z; x[3] = a.w;
    x[4] = b.x; x[5] = b.y; x[6] = b.z; x[7] = b.w;
}

__device__ __forceinline__ void store_elem(const uint32_t* x, uint32_t* base, int layout, uint64_t n,
                                           uint64_t i, uint32_t w, uint32_t k) {
    uint4* p0;
    uint4* p1;
    if (layout == kLayoutAos) {
        p0 = reinterpret_cast<uint4*>(base + (i * w + k) * 8);
        p1 = p0 + 1;
    } else {
        p0 = reinterpret_cast<uint4*>(base + ((uint64_t)(k * 2) * n + i) * 4);
        p1 = reinterpret_cast<uint4*>(base + ((uint64_t)(k * 2 + 1) * n + i) * 4);
    }
    *p0 = make_uint4(x[0], x[1], x[2], x[3]);
    *p1 = make_uint4(x[4], x[5], x[6], x[7]);
}

// ------------------------------------------------------------------------------------------------
// The sponge kernel.  VARIANT selects the permutation implementation (0 = dense/reference order).
// ------------------------------------------------------------------------------------------------
template <class F, int T, int ALPHA, int VARIANT>
__device__ __forceinline__ void permute(uint32_t (&st)[T][8], const PermDesc& d) {
    static_assert(VARIANT < 6, "variants 6 and 7 are called with their warp / tile context (permute_tc, permute_tc5)");
    if constexpr (VARIANT == 5) {
        permute_fp52<F, T, ALPHA>(st, d);       // the host only selects this kernel when d.fp_ok
    } else if constexpr (T > 5) {
        permute_rolled<F, T, ALPHA>(st, d);
    } else if constexpr (VARIANT == 4) {
        permute_opt_ctab<F, T, ALPHA, true>(st, d);    // lane-update tables only (d.ctab_ok == 2)
    } else if constexpr (VARIANT == 3) {
        permute_opt_ctab<F, T, ALPHA>(st, d);   // the host only selects this kernel when d.ctab_ok
    } else if constexpr (VARIANT == 2) {
        permute_opt_compact<F, T, ALPHA>(st, d);
    } else if constexpr (VARIANT == 1) {
        permute_opt<F, T, ALPHA>(st, d);     // the host only selects this kernel when d.opt_ok
    } else {
        permute_dense<F, T, ALPHA>(st, d);
    }
}

// Resident 128-thread blocks per SM the register allocation is steered towards (the multiplier pipe
// needs >= 4 warps per scheduler to stay busy; see profiles/).  Overridable for tuning runs.
// Measured on B200: with round constants in uniform registers the t=3 kernels need only 72 vector
// registers, and 7 resident blocks (7 warps per scheduler) are best: c2 147.1 M perm/s vs 145.0 M at 5,
// 146.2 M at 8 (spills); c5 159.2 M vs 156.0 M.  t=5 is best when ptxas is left alone.
#ifndef PSD_MIN_BLOCKS
#define PSD_MIN_BLOCKS(T) ((T) <= 3 ? 7 : (T) == 5 ? 5 : 0)   /* 0 = no hint; t=5: 83.8 M vs 83.0 M hash/s */
#endif
// Variant 5 keeps the state in 52-bit limb pairs next to the integer S-box temporaries, and its partial
// rounds hold an S-box and two lane updates in one block: 128 registers per thread.  It may be launched
// with up to 512 threads per block (one such block fills an SM, see sponge_cfg in api.cu); every other
// variant runs 128-thread blocks.  Measured on B200 (c2, 128-thread blocks): 163.1 M perm/s at 128
// registers, 147.3 M when capped at 96 (spills); not occupancy-limited -- the form with rolled phases
// ran 162 M at 5, 6 and 7 resident blocks alike.
#define PSD_KERNEL_MAX_BLOCK(VARIANT) ((VARIANT) == 5 ? 512 : 128)
#define PSD_KERNEL_BLOCKS_AT_MAX(T, VARIANT) ((VARIANT) == 5 ? 1 : PSD_MIN_BLOCKS(T))
// Shared-memory form (POOL == 1): every block carries its own copy of the constant image, so blocks are
// large -- variant 5 as above (512 threads, one block per SM), the integer variants 448 threads with two
// blocks per SM (the 28 warps per SM that seven 128-thread blocks give the constant-bank form).
// Variant 6: block size is a build-time choice (PSD_TC_BLOCK, default 512): larger blocks mean more warps per
// scheduler to hide the S-box's dependency chains, at fewer registers per thread.
#ifndef PSD_TC_BLOCK
#define PSD_TC_BLOCK 512
#endif
// Variant 7: whole 128-sponge tiles; a tile owns 32 T of the SM's 512 TMEM columns, so five tiles (96 registers per thread)
// up to T = 3, four (128 registers) at T = 4, three (168 registers) at T = 5.
#ifndef PSD_TC5_BLOCK
#define PSD_TC5_BLOCK(T) ((T) <= 3 ? 640 : (T) == 4 ? 512 : 384)
#endif
#define PSD_SMEM_BLOCK(T, VARIANT) ((VARIANT) == 6 ? PSD_TC_BLOCK : (VARIANT) == 7 ? PSD_TC5_BLOCK(T) : (VARIANT) == 5 ? 512 : 448)
#define PSD_SMEM_BLOCKS(VARIANT) ((VARIANT) >= 5 ? 1 : 2)
#if defined(PSD_POOL_SHARED) && PSD_POOL_SHARED
#define PSD_POOL_KIND 1
#define PSD_LAUNCH_BOUNDS(T, VARIANT) __launch_bounds__(PSD_SMEM_BLOCK(T, VARIANT), PSD_SMEM_BLOCKS(VARIANT))
#define PSD_MAX_BLOCK(VARIANT) PSD_SMEM_BLOCK(3, VARIANT)
#define PSD_MAX_BLOCK_T(T, VARIANT) PSD_SMEM_BLOCK(T, VARIANT)
#else
#define PSD_POOL_KIND 0
#define PSD_LAUNCH_BOUNDS(T, VARIANT) __launch_bounds__(PSD_KERNEL_MAX_BLOCK(VARIANT), PSD_KERNEL_BLOCKS_AT_MAX(T, VARIANT))
#define PSD_MAX_BLOCK(VARIANT) PSD_KERNEL_MAX_BLOCK(VARIANT)
#endif
// POOL is part of the kernel's name only so that the constant-bank and the shared-memory build of the same
// (field, T, ALPHA, VARIANT) are different symbols; the code difference comes from how c_pool is declared
// in the translation unit (poseidon_core.cuh).
//
// RAGGED: every sponge absorbs its own number of elements (a.in_offsets), as the reference's callers do when
// they absorb lists of differing lengths into different sponges (src/poseidon/tests.rs:45-69,
// absorb.rs:82-101).  The sequencer's variables are per thread already; with differing lengths the threads of
// a warp simply leave the absorb loop at different times (threads that need no permutation in an iteration
// idle through it).  A separate instantiation, because inside divergent control flow ptxas keeps the round
// constants out of the uniform datapath -- the lock-step kernels must not pay for that.
template <class F, int T, int ALPHA, int VARIANT, int POOL = PSD_POOL_KIND, bool RAGGED = false>
__global__ void PSD_LAUNCH_BOUNDS(T, VARIANT) sponge_kernel(const __grid_constant__ SpongeArgs a) {
    static_assert(POOL == PSD_POOL_KIND, "POOL must match the translation unit's pool kind");
    const PermDesc& d = a.d;
    const uint32_t rate = d.rate, cap = d.capacity;
    bool bad = false;

#if PSD_POOL_KIND == 1
    {   // stage the context's constant image: global (L2-resident after the first blocks) -> shared
        const uint4* src = reinterpret_cast<const uint4*>(a.pool_img);
        uint4* dst = reinterpret_cast<uint4*>(c_pool);
        const uint32_t n16 = a.pool_words >> 2;
        for (uint32_t q = threadIdx.x; q < n16; q += blockDim.x) dst[q] = __ldg(src + q);
        __syncthreads();
    }
#endif

    // One sponge per thread and an early exit -- deliberately NOT a grid-stride loop: inside a
    // (potentially divergent) per-thread loop ptxas stops using the uniform datapath, and every
    // round-constant read turns from LDCU + uniform-register operand into LEA + LDC into a vector
    // register (measured: +6 % instructions on the compact kernel, +10 % on the digit-table one).
    // Variant 6 runs warp-collective instructions (mma.sync, ldmatrix) inside the permutation, so the threads of
    // the last, partly filled warp stay alive: they recompute the last sponge and store nothing.
    const uint64_t i_raw = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const bool live = i_raw < a.n;
    if (VARIANT < 6 && !live) return;
    const uint64_t i = live ? i_raw : a.n - 1;
#if PSD_POOL_KIND == 1
    TcWarp tcw;
    if (VARIANT == 6) {
        // per-warp scratch behind the staged constant image (128-byte aligned)
        unsigned char* base = reinterpret_cast<unsigned char*>(c_pool) + ((a.pool_words * 4u + 127u) & ~127u);
        tcw.scratch = base + (threadIdx.x >> 5) * kTcWarpScratch;
        tcw.tab = a.tc_tab;
        tcw.lane = threadIdx.x & 31;
        tcw.a_use = tcw.c_use = 0;
        // The warps of a block start together and would stay in phase: all in the S-box (integer multiplier busy,
        // tensor cores and shared memory idle), then all in the linear layer (the reverse).  Starting the warps that
        // share a scheduler (w, w+4, w+8, w+12) a fraction of a round apart lets the two halves overlap.
        if (a.tc_stagger) {
            const long long t0 = clock64();
            const long long wait = (long long)((threadIdx.x >> 7) & 3) * a.tc_stagger;
            while (clock64() - t0 < wait) {
            }
        }
    }
    // Variant 7: the block is blockDim.x / 128 tiles of 128 sponges.  Shared memory behind the staged constant image:
    // [M00, ident, reduction tables][A slots per tile][U: matrices / per-tile staging], then the tiles' mbarriers
    // (poseidon_tc5.cuh).
    Tc5Tile tile;
    if (VARIANT == 7) {
        unsigned char* base = reinterpret_cast<unsigned char*>(c_pool) + ((a.pool_words * 4u + 1023u) & ~1023u);
        // the tile index through a shuffle: warp-uniform to the compiler, so that everything derived from it (operand
        // addresses, descriptors, the issuing branch) stays on the uniform datapath
        const uint32_t n_tiles = blockDim.x >> 7, tl = __shfl_sync(0xffffffffu, threadIdx.x >> 7, 0);
        unsigned char* a0 = base + tc5_small_bytes();
        unsigned char* u0 = a0 + n_tiles * tc5_a_bytes(T);
        unsigned char* bars = u0 + tc5_u_bytes(T, n_tiles);
        uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 120);
        tile.row = threadIdx.x & 127;
        tile.leader_warp = __shfl_sync(0xffffffffu, (threadIdx.x >> 5) & 3, 0) == 0;
        tile.bar_id = 1 + (int)tl;
        tile.tab = a.tc_tab;
        tile.a_ptr = a0 + tl * tc5_a_bytes(T);
        tile.a_q = (uint32_t)__cvta_generic_to_shared(tile.a_ptr) >> 4;
        tile.small_q = (uint32_t)__cvta_generic_to_shared(base) >> 4;
        tile.u_u32 = (uint32_t)__cvta_generic_to_shared(u0);
        tile.stage_u32 = tile.u_u32 + tl * tc5_stage_bytes(T);
        tile.mbar_u32[0] = (uint32_t)__cvta_generic_to_shared(bars + tl * 16);
        tile.mbar_u32[1] = tile.mbar_u32[0] + 8;
        tile.parity[0] = tile.parity[1] = 0;
        if (tile.row == 0) {
            asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(tile.mbar_u32[0]));
            asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(tile.mbar_u32[1]));
            asm volatile("fence.mbarrier_init.release.cluster;");
        }
        // the block's small tables: M00 (2 KiB), ident (1 KiB), the reduction table (2 KiB)
        {
            const uint4* g = reinterpret_cast<const uint4*>(a.tc_tab);
            uint4* sdst = reinterpret_cast<uint4*>(base);
            for (uint32_t q = threadIdx.x; q < 128; q += blockDim.x) sdst[q] = __ldg(g + a.tc.m00_off / 16 + q);
            for (uint32_t q = threadIdx.x; q < 64; q += blockDim.x) sdst[128 + q] = __ldg(g + a.tc.ident_off / 16 + q);
            for (uint32_t q = threadIdx.x; q < 128; q += blockDim.x) sdst[192 + q] = __ldg(g + a.tc.red_off / 16 + q);
        }
        // tensor memory: 32 T columns per tile (a power of two >= 32 in all)
        const uint32_t cols = tc5_tmem_cols(T, n_tiles);
        if (threadIdx.x < 32) {
            asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"((uint32_t)__cvta_generic_to_shared(tmem_slot)), "r"(cols));
            asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
        }
        asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
        // the dense matrices into U; ends with a block-wide barrier behind a proxy fence (covers the small tables, the
        // mbarriers and the TMEM address as well)
        tc5_load_matrices<T>(tile, a.tc.mat_off, (tc5_both_matrices(T) && a.d.partial != 0) ? a.tc.pre_off : 0);
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        tile.tmem = __shfl_sync(0xffffffffu, *tmem_slot, 0) + tl * 32 * T;
    }
#endif
    {
        uint32_t st[T][8];
        if (a.state_in != nullptr) {
#pragma unroll
            for (int l = 0; l < T; l++) {
                load_elem(st[l], a.state_in, a.state_in_layout, a.n, i, T, l);
                if (a.state_in_repr == kReprCanonical) {
                    bad |= !is_canonical<F>(st[l]);
                    to_mont<F>(st[l], d);
                }
            }
        } else {
#pragma unroll
            for (int l = 0; l < T; l++)
#pragma unroll
                for (int k = 0; k < 8; k++) st[l][k] = 0;
        }

        // Warp-uniform sequencer with a single permutation site.
        int stage = 0;
        int mode = a.mode;
        uint32_t idx = a.index;
        uint32_t rem_raw = a.n_raw;
        uint32_t n_abs = a.n_absorb;
        const uint32_t* in_base = a.in;
        if (RAGGED) {
            const uint64_t lo = a.in_offsets[i], hi = a.in_offsets[i + 1];
            n_abs = (uint32_t)(hi - lo);
            in_base = a.in + (lo - a.in_offset_base) * 8;
        }
        uint32_t rem_in = n_abs, k_in = 0;
        uint32_t rem_out = a.n_squeeze, k_out = 0;
        for (;;) {
            bool perm = false;
            uint32_t add_cnt = 0, emit_cnt = 0, lane0 = 0;
            switch (stage) {
                case 0:   // raw permutations
                    if (rem_raw) {
                        rem_raw--;
                        perm = true;
                    } else {
                        stage = 1;
                    }
                    break;
                case 1:   // CryptographicSponge::absorb entry, mod.rs:232-254
                    if (n_abs == 0) {
                        stage = 3;
                        break;
                    }
                    if (mode == kModeAbsorbing) {
                        if (idx == rate) {
                            perm = true;
                            idx = 0;
                        }
                    } else {
                        perm = true;
                        idx = 0;
                    }
                    mode = kModeAbsorbing;
                    stage = 2;
                    break;
                case 2:   // absorb_internal, mod.rs:121-150
                    lane0 = cap + idx;
                    if (idx + rem_in <= rate) {
                        add_cnt = rem_in;
                        idx += rem_in;
                        stage = 3;
                    } else {
                        add_cnt = rate - idx;
                        idx = 0;
                        perm = true;
                    }
                    break;
                case 3:   // squeeze_native_field_elements entry, mod.rs:320-342
                    if (!a.do_squeeze) {
                        stage = 5;
                        break;
                    }
                    if (mode == kModeAbsorbing || idx == rate) {
                        perm = true;
                        idx = 0;
                    }
                    mode = kModeSqueezing;
                    stage = 4;
                    break;
                case 4:   // squeeze_internal, mod.rs:153-182
                    lane0 = cap + idx;
                    if (idx + rem_out <= rate) {
                        emit_cnt = rem_out;
                        idx += rem_out;
                        stage = 5;
                    } else {
                        emit_cnt = rate - idx;
                        perm = (rem_out != rate);   // tested before the decrement, mod.rs:175
                        idx = 0;
                    }
                    break;
                default:
                    break;
            }

            if (add_cnt) {
#pragma unroll
                for (int l = 0; l < T; l++) {
                    const uint32_t j = (uint32_t)l - lane0;
                    if ((uint32_t)l >= lane0 && j < add_cnt) {
                        uint32_t e[8];
                        if (RAGGED) load_elem(e, in_base, kLayoutAos, 1, 0, 0, k_in + j);
                        else load_elem(e, a.in, a.in_layout, a.n, i, a.n_absorb, k_in + j);
                        if (a.in_repr == kReprCanonical) {
                            bad |= !is_canonical<F>(e);
                            to_mont<F>(e, d);
                        }
                        fp_add<F>(st[l], st[l], e);
                    }
                }
                k_in += add_cnt;
                rem_in -= add_cnt;
            }
            if (emit_cnt) {
#pragma unroll
                for (int l = 0; l < T; l++) {
                    const uint32_t j = (uint32_t)l - lane0;
                    if ((uint32_t)l >= lane0 && j < emit_cnt) {
                        uint32_t e[8];
#pragma unroll
                        for (int k = 0; k < 8; k++) e[k] = st[l][k];
                        if (VARIANT >= 6 && !live) {
                            // a padding thread of the last warp: nothing to deliver
                        } else if (a.out_bytes != nullptr) {
                            // squeeze_bytes, mod.rs:256-270: first 31 LE bytes of the canonical value
                            from_mont<F>(e);
                            const uint32_t usable = (F::kBits - 1) / 8;
                            const uint32_t off = (k_out + j) * usable;
                            uint8_t* dst = a.out_bytes + i * (uint64_t)a.n_bytes + off;
                            for (uint32_t b = 0; b < usable && off + b < a.n_bytes; b++)
                                dst[b] = (uint8_t)(e[b >> 2] >> ((b & 3) * 8));
                        } else {
                            if (a.out_repr == kReprCanonical) from_mont<F>(e);
                            store_elem(e, a.out, a.out_layout, a.n, i, a.n_squeeze, k_out + j);
                        }
                    }
                }
                k_out += emit_cnt;
                rem_out -= emit_cnt;
            }
#if PSD_POOL_KIND == 1
            if constexpr (VARIANT == 7 && RAGGED) {
                // Per-sponge lengths under a kernel whose permutation is a collective of the whole block (tile barriers,
                // the block-wide switch of U): the block permutes as long as ANY of its sponges needs to.  A sponge that
                // does not need this permutation runs it all the same and puts its state back afterwards.
                const bool done = !perm && stage == 5;
                if (__syncthreads_or(perm ? 1 : 0)) {
                    uint32_t keep[T][8];
                    if (!perm) {
#pragma unroll
                        for (int l = 0; l < T; l++)
#pragma unroll
                            for (int k = 0; k < 8; k++) keep[l][k] = st[l][k];
                    }
                    permute_tc5<F, T, ALPHA>(st, d, a.tc, tile);
                    if (!perm) {
#pragma unroll
                        for (int l = 0; l < T; l++)
#pragma unroll
                            for (int k = 0; k < 8; k++) st[l][k] = keep[l][k];
                    }
                } else if (__syncthreads_and(done ? 1 : 0)) {
                    break;
                }
                continue;
            }
#endif
            if (perm) {
#if PSD_POOL_KIND == 1
                if constexpr (VARIANT == 6) permute_tc<F, T, ALPHA>(st, d, a.tc, tcw);
                else if constexpr (VARIANT == 7) permute_tc5<F, T, ALPHA>(st, d, a.tc, tile);
                else
#endif
                    permute<F, T, ALPHA, VARIANT>(st, d);
            } else if (stage == 5) {
                break;
            }
        }

        if (a.state_out != nullptr && (VARIANT < 6 || live)) {
#pragma unroll
            for (int l = 0; l < T; l++) {
                uint32_t e[8];
#pragma unroll
                for (int k = 0; k < 8; k++) e[k] = st[l][k];
                if (a.state_out_repr == kReprCanonical) from_mont<F>(e);
                store_elem(e, a.state_out, a.state_out_layout, a.n, i, T, l);
            }
        }
    }
#if PSD_POOL_KIND == 1
    if (VARIANT == 7) {
        asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
        __syncthreads();
        if (threadIdx.x < 32) {
            const uint32_t cols = tc5_tmem_cols(T, blockDim.x >> 7);
            asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tile.tmem), "r"(cols));   // warp 0: tile 0
        }
    }
#endif
    if (bad && a.err && live) atomicOr(a.err, 1);
}

}  // namespace psd
