// Device-side Poseidon permutation and duplex-sponge step, one sponge per thread, state in registers.
//
// Follows /root/reference/src/poseidon/mod.rs:
//   apply_ark 76-80, apply_s_box 63-74, apply_mds 82-93, permute 95-118,
//   absorb_internal 121-150, squeeze_internal 153-182, absorb 232-254,
//   squeeze_native_field_elements 320-342.
// All round constants live in the constant bank (c_pool) in Montgomery form and are read at
// warp-uniform addresses, so they reach the multiplier as uniform-register operands.
#pragma once
#include "fp256.cuh"

namespace psd {

// 62 KiB of the 64 KiB constant bank, shared by all live contexts of a device (slot allocator on
// the host side, see api.cu).
constexpr uint32_t kConstPoolWords = 15872;
// Largest constant image a kernel of the shared-memory form accepts (one block per SM may use up to
// 227 KiB; the t = 5 variant-5 image is the largest in use at about 100 KiB).
constexpr uint32_t kSmemPoolMaxWords = 56 * 1024;

// Where the kernels of a translation unit read their constants from.  Default: the unit's slice of the
// constant bank (uniform loads, operands in uniform registers).  With -DPSD_POOL_SHARED=1 the same code
// is compiled against a dynamic shared-memory array that every block fills from the context's image in
// global memory before its first permutation (sponge_kernel's prologue): no 62 KiB limit and no
// competition between parameter sets, at the price of broadcast LDS reads into vector registers.
#if defined(__CUDACC__)
#if defined(PSD_POOL_SHARED) && PSD_POOL_SHARED
extern __shared__ __align__(16) uint32_t c_pool[];
#else
__constant__ __align__(16) uint32_t c_pool[kConstPoolWords];
#endif
#else
alignas(16) static uint32_t c_pool[kSmemPoolMaxWords];
#endif

// One field element (8 words) from the pool.  Offsets are multiples of 8 words, so this is two
// 128-bit uniform loads (LDCU.128) with one address computation instead of eight 32-bit ones.
PSD_DEV void cload8(uint32_t* dst, uint32_t off) {
#if defined(__CUDACC__)
    // indexing in 16-byte units makes the alignment provable (off is a runtime multiple of 8 words)
    const uint4* q = reinterpret_cast<const uint4*>(c_pool) + (off >> 2);
    const uint4 lo = q[0];
    const uint4 hi = q[1];
    dst[0] = lo.x; dst[1] = lo.y; dst[2] = lo.z; dst[3] = lo.w;
    dst[4] = hi.x; dst[5] = hi.y; dst[6] = hi.z; dst[7] = hi.w;
#else
    for (int k = 0; k < 8; k++) dst[k] = c_pool[off + k];
#endif
}

// What a kernel needs to know about one parameter set.  Offsets are in 32-bit words into c_pool.
struct PermDesc {
    uint32_t ark_off;     // [rounds][T][8]
    uint32_t mds_off;     // [T][T][8]   (row-major: mds[i][j] multiplies state[j] for output i)
    uint32_t one_off;     // [8]  Montgomery 1 (for the generic pow ladder)
    uint32_t r2_off;      // [8]  R^2 mod p   (canonical -> Montgomery)
    uint32_t half_full;   // full_rounds / 2          (poseidon/mod.rs:96)
    uint32_t partial;     // partial_rounds
    uint32_t total;       // full_rounds + partial_rounds
    uint32_t rate;
    uint32_t capacity;
    uint32_t alpha_lo, alpha_hi;
    // optimised form (see opt_constants.cpp); valid when opt_ok != 0
    uint32_t opt_ok;
    uint32_t pre_mds_off;   // [T][T][8]  matrix of the last full round before the partial rounds
    uint32_t sparse_off;    // [partial][2T][8]: k, M00, w_1..w_{T-1}, v_1..v_{T-1} per partial round
    uint32_t post_ark_off;  // [T][8]     round constants of the first full round after the partial rounds
    // digit tables for constant products (fp256.cuh const_dot); valid when ctab_ok != 0.  One table =
    // 4 entries x 8 words.
    uint32_t ctab_ok;       // 0 = none, 1 = all tables, 2 = lane-update tables only (ct_sparse_off: [partial][T-1][4][8])
    uint32_t ct_m00_off;    // [4][8]                table of mds[0][0]
    uint32_t ct_mds_off;    // [T][T][4][8]          tables of the dense matrix
    uint32_t ct_pre_off;    // [T][T][4][8]          tables of M_pre
    uint32_t ct_sparse_off; // [partial][2(T-1)][4][8]: w_1..w_{T-1}, v_1..v_{T-1} per partial round
    // FP64-pipe constant products (fp52.cuh / poseidon_fp52.cuh, VARIANT 5); valid when fp_ok != 0.
    // Offsets in 8-BYTE units into the pool of the unit that holds the variant-5 kernels.
    uint32_t fp_ok;
    uint32_t fp_consts_off; // P52[5], NP52, SHIFT, 2^104
    uint32_t fp_mat_off;    // [T][T][5][5] doubles: tables of the dense matrix (row-major; [0][0] doubles as M00)
    uint32_t fp_pre_off;    // same shape: M_pre
    uint32_t fp_last_off;   // same shape: the matrix of the last partial round (sparse row + lane updates + rescale)
    uint32_t fp_k_off;      // additive constants, 6 int64 per row, in execution order
    uint32_t fp_part_off;   // [partial][2(T-1)][5][5] doubles: w_1.., v_1.. tables per partial round
};

template <class F, int ALPHA>
PSD_DEV void sbox(uint32_t* x, const PermDesc& d) {
    if (ALPHA == 5) {
        fp_pow5<F>(x, x);
    } else if (ALPHA == 17) {
        uint32_t y[8];
        fp_sqr<F>(y, x);
        fp_sqr<F>(y, y);
        fp_sqr<F>(y, y);
        fp_sqr<F>(y, y);
        mont_mul<F>(x, y, x);
    } else {
        const uint64_t alpha = ((uint64_t)d.alpha_hi << 32) | d.alpha_lo;
        uint32_t one[8];
        cload8(one, d.one_off);
        fp_pow_u64<F>(x, x, alpha, one);
    }
}

// One round exactly as the reference orders it: ark -> S-box -> dense MDS.
template <class F, int T, int ALPHA>
PSD_DEV void round_dense(uint32_t (&st)[T][8], const PermDesc& d, uint32_t r, bool full) {
    // apply_ark
#pragma unroll
    for (int i = 0; i < T; i++) {
        uint32_t c[8];
        cload8(c, d.ark_off + (r * T + i) * 8);
        fp_add<F>(st[i], st[i], c);
    }
    // apply_s_box
    if (full) {
#pragma unroll
        for (int i = 0; i < T; i++) sbox<F, ALPHA>(st[i], d);
    } else {
        sbox<F, ALPHA>(st[0], d);
    }
    // apply_mds: new[i] = sum_j st[j] * mds[i][j]
    uint32_t nw[T][8];
#pragma unroll
    for (int i = 0; i < T; i++) {
#pragma unroll
        for (int j = 0; j < T; j++) {
            uint32_t m[8], prod[8];
            cload8(m, d.mds_off + (i * T + j) * 8);
            mont_mul<F>(prod, st[j], m);
            if (j == 0) {
#pragma unroll
                for (int k = 0; k < 8; k++) nw[i][k] = prod[k];
            } else {
                fp_add<F>(nw[i], nw[i], prod);
            }
        }
    }
#pragma unroll
    for (int i = 0; i < T; i++)
#pragma unroll
        for (int k = 0; k < 8; k++) st[i][k] = nw[i][k];
}

// permute, poseidon/mod.rs:95-118
template <class F, int T, int ALPHA>
PSD_DEV void permute_dense(uint32_t (&st)[T][8], const PermDesc& d) {
#if defined(__CUDACC__)
#pragma unroll 1
#endif
    for (uint32_t r = 0; r < d.total; r++) {
        const bool full = (r < d.half_full) || (r >= d.half_full + d.partial);
        round_dense<F, T, ALPHA>(st, d, r, full);
    }
}

// ------------------------------------------------------------------------------------------------
// Optimised permutation: same function, fewer multiplier-pipe cycles.
//  * partial rounds in sparse form: the MDS matrix is factored (host, once) so that a partial round
//    is   s0 <- (s0 + k)^alpha ;  s0' = M00*s0 + sum_j w_j*s_j ;  s_j' = s_j + v_j*s0   (2T-1 products
//    instead of T^2, one round constant instead of T);
//  * every matrix row is a lazily reduced dot product (mont_dot) where the field allows it.
// Results are identical to permute_dense: the values are the same field elements, fully reduced.
// ------------------------------------------------------------------------------------------------
// part = sum_{j<n} a[j]*cf[j] (n <= G <= 4 known at compile time after unrolling)
template <class F, int G>
PSD_DEV void dot_group(uint32_t* part, const uint32_t (*a)[8], const uint32_t (*cf)[8], int n) {
    if (n == 1) {
        mont_mul<F>(part, a[0], cf[0]);
    } else if (DotLimit<F>::kWide) {
        if (n == 2) mont_dot_w<F, 2>(part, a, cf);
        else if (n == 3) mont_dot_w<F, (G >= 3 ? 3 : 1)>(part, a, cf);
        else mont_dot_w<F, (G >= 4 ? 4 : 1)>(part, a, cf);
    } else {
        if (n == 2) mont_dot<F, 2>(part, a, cf);
        else if (n == 3) mont_dot<F, (G >= 3 ? 3 : 1)>(part, a, cf);
        else mont_dot<F, (G >= 4 ? 4 : 1)>(part, a, cf);
    }
}

template <class F, int T>
PSD_DEV void row_dot(uint32_t* r, const uint32_t (*st)[8], uint32_t coeff_off) {
    // r = sum_{j<T} st[j] * c_pool[coeff_off + 8j ...], in groups of at most DotLimit<F>::kMax terms
    if (T == 5 && F::kBits == 254) {
        // BN254, five terms: one lazily reduced group on the wide window (6p > 2^256 rules out the
        // 9-limb window; the result is < 1.95p, within reduce_3p's range) -- one reduction instead of two
        uint32_t cf[T][8];
#pragma unroll
        for (int j = 0; j < T; j++)
            cload8(cf[j], coeff_off + j * 8);
        mont_dot_w<F, T>(r, st, cf);
        return;
    }
    constexpr int G = DotLimit<F>::kMax;
    bool first = true;
#pragma unroll
    for (int j0 = 0; j0 < T; j0 += G) {
        uint32_t part[8];
        uint32_t cf[G][8];
#pragma unroll
        for (int j = 0; j < G; j++) {
            if (j0 + j < T) {
                cload8(cf[j], coeff_off + (j0 + j) * 8);
            } else {
#pragma unroll
                for (int k = 0; k < 8; k++) cf[j][k] = 0u;
            }
        }
        const int n = (T - j0 < G) ? (T - j0) : G;
        dot_group<F, G>(part, st + j0, cf, n);
        if (first) {
#pragma unroll
            for (int k = 0; k < 8; k++) r[k] = part[k];
            first = false;
        } else {
            fp_add<F>(r, r, part);
        }
    }
}

template <class F, int T, int ALPHA>
PSD_DEV void full_round_opt(uint32_t (&st)[T][8], const PermDesc& d, uint32_t ark_off, uint32_t mds_off) {
#pragma unroll
    for (int i = 0; i < T; i++) {
        uint32_t c[8];
        cload8(c, ark_off + i * 8);
        fp_add<F>(st[i], st[i], c);
        sbox<F, ALPHA>(st[i], d);
    }
    uint32_t nw[T][8];
#pragma unroll
    for (int i = 0; i < T; i++) row_dot<F, T>(nw[i], st, mds_off + i * T * 8);
#pragma unroll
    for (int i = 0; i < T; i++)
#pragma unroll
        for (int k = 0; k < 8; k++) st[i][k] = nw[i][k];
}

template <class F, int T, int ALPHA>
PSD_DEV void partial_round_opt(uint32_t (&st)[T][8], const PermDesc& d, uint32_t off) {
    // off -> [k][M00][w_1..w_{T-1}][v_1..v_{T-1}]
    {
        uint32_t c[8];
        cload8(c, off);
        fp_add<F>(st[0], st[0], c);
        sbox<F, ALPHA>(st[0], d);
    }
    // s0' = M00*s0 + sum w_j s_j  : coefficients (M00, w_1, ...) stored contiguously after k
    uint32_t n0[8];
    row_dot<F, T>(n0, st, off + 8);
    // s_j' = s_j + v_j * s0
#pragma unroll
    for (int j = 1; j < T; j++) {
        uint32_t v[8], prod[8];
        cload8(v, off + (T + j) * 8);
        mont_mul<F>(prod, st[0], v);
        fp_add<F>(st[j], st[j], prod);
    }
#pragma unroll
    for (int k = 0; k < 8; k++) st[0][k] = n0[k];
}

template <class F, int T, int ALPHA>
PSD_DEV void permute_opt(uint32_t (&st)[T][8], const PermDesc& d) {
    const uint32_t first_partial = d.half_full, first_post = d.half_full + d.partial;
#if defined(__CUDACC__)
#pragma unroll 1
#endif
    for (uint32_t r = 0; r < d.total; r++) {
        if (r < first_partial || r >= first_post) {
            const uint32_t ark = (r == first_post && d.partial) ? d.post_ark_off : d.ark_off + r * T * 8;
            const uint32_t mds = (r + 1 == first_partial && d.partial) ? d.pre_mds_off : d.mds_off;
            full_round_opt<F, T, ALPHA>(st, d, ark, mds);
        } else {
            partial_round_opt<F, T, ALPHA>(st, d, d.sparse_off + (r - first_partial) * (2 * T) * 8);
        }
    }
}

// ------------------------------------------------------------------------------------------------
// Compact optimised permutation (VARIANT 2): the same arithmetic as permute_opt, but every round is
// run as three short rolled phases so that only ONE copy each of the S-box, the row dot product and
// the lane multiply-add exists in the hot loop (about a third of the code of permute_opt; ncu showed
// 21 % of warp stalls of permute_opt as no_instruction, i.e. instruction fetch).  Lanes are rotated
// through fixed register names with moves instead of being indexed.
//   phase A: nA x { s0 += c_i; s0 <- s0^alpha; [rotate lanes] }        nA = T (full) | 1 (partial)
//   phase B: nB x { t = <coeff row i, s>; shift t into n }             nB = T (full) | 1 (partial)
//   phase C: partial only, (T-1) x { s1 += v_j * s0; rotate lanes 1..T-1 }
// ------------------------------------------------------------------------------------------------
// LAZY (BN254 Fr only, p = 0.189 * 2^256): the three conditional subtractions inside x^alpha are
// dropped and the result is left in [0, 1.25 p).  With x < p on entry: x^2 < (1 + 0.189) p, and a
// squaring of a < 1.33 p stays below (1 + 1.33^2 * 0.189) p < 1.34 p, inside mont_sqr's operand bound
// 2a + p < 2^256; the closing product has a = y < 1.34 p (a + p < 2^256) and b = x < p, giving
// < (1 + 1.34 * 0.189) p < 1.26 p.  Only callers whose consumers take any 256-bit value may ask for
// it: const_dot reads the value digit by digit and its bounds do not depend on the operand.
template <class F, int ALPHA, bool LAZY = false>
PSD_DEV void sbox_compact(uint32_t* x, const PermDesc& d) {
    static_assert(!LAZY || (F::kBits == 254 && (ALPHA == 5 || ALPHA == 17)), "lazy S-box: BN254, alpha 5/17 only");
    if (ALPHA == 5 || ALPHA == 17) {
        uint32_t y[8];
#pragma unroll
        for (int k = 0; k < 8; k++) y[k] = x[k];
#if defined(__CUDACC__)
#pragma unroll 1
#endif
        for (int q = 0; q < (ALPHA == 5 ? 2 : 4); q++) fp_sqr<F, !LAZY>(y, y);
        mont_mul<F, !LAZY>(x, y, x);
    } else {
        sbox<F, ALPHA>(x, d);
    }
}

template <class F, int T, int ALPHA>
PSD_DEV void permute_opt_compact(uint32_t (&st)[T][8], const PermDesc& d) {
    const uint32_t first_partial = d.half_full, first_post = d.half_full + d.partial;
#if defined(__CUDACC__)
#pragma unroll 1
#endif
    for (uint32_t r = 0; r < d.total; r++) {
        const bool full = (r < first_partial || r >= first_post);
        const uint32_t sp = d.sparse_off + (r - first_partial) * (2 * T) * 8;   // only meaningful when !full
        const uint32_t ark = full ? ((r == first_post && d.partial) ? d.post_ark_off : d.ark_off + r * T * 8) : sp;
        const uint32_t coeff = full ? ((r + 1 == first_partial && d.partial) ? d.pre_mds_off : d.mds_off) : sp + 8;
        const uint32_t n_iter = full ? T : 1;
        // phase A
#if defined(__CUDACC__)
#pragma unroll 1
#endif
        for (uint32_t i = 0; i < n_iter; i++) {
            uint32_t c[8];
            cload8(c, ark + i * 8);
            fp_add<F>(st[0], st[0], c);
            sbox_compact<F, ALPHA>(st[0], d);
            if (full) {   // rotate left: the next lane moves into position 0
#pragma unroll
                for (int k = 0; k < 8; k++) {
                    const uint32_t t0 = st[0][k];
#pragma unroll
                    for (int l = 0; l + 1 < T; l++) st[l][k] = st[l + 1][k];
                    st[T - 1][k] = t0;
                }
            }
        }
        // phase B
        uint32_t nw[T][8];
#pragma unroll
        for (int l = 0; l < T; l++)
#pragma unroll
            for (int k = 0; k < 8; k++) nw[l][k] = 0;
#if defined(__CUDACC__)
#pragma unroll 1
#endif
        for (uint32_t i = 0; i < n_iter; i++) {
            uint32_t t[8];
            row_dot<F, T>(t, st, coeff + i * T * 8);
#pragma unroll
            for (int k = 0; k < 8; k++) {
#pragma unroll
                for (int l = 0; l + 1 < T; l++) nw[l][k] = nw[l + 1][k];
                nw[T - 1][k] = t[k];
            }
        }
        if (full) {
#pragma unroll
            for (int l = 0; l < T; l++)
#pragma unroll
                for (int k = 0; k < 8; k++) st[l][k] = nw[l][k];
        } else {
            // phase C: s_j += v_j * s0 for j = 1..T-1, lanes 1..T-1 rotated through position 1
#if defined(__CUDACC__)
#pragma unroll 1
#endif
            for (uint32_t j = 0; j + 1 < T; j++) {
                uint32_t v[8], prod[8];
                cload8(v, sp + (T + 1 + j) * 8);
                mont_mul<F>(prod, st[0], v);
                fp_add<F>(st[1], st[1], prod);
#pragma unroll
                for (int k = 0; k < 8; k++) {
                    const uint32_t t1 = st[1][k];
#pragma unroll
                    for (int l = 1; l + 1 < T; l++) st[l][k] = st[l + 1][k];
                    st[T - 1][k] = t1;
                }
            }
#pragma unroll
            for (int k = 0; k < 8; k++) st[0][k] = nw[T - 1][k];
        }
    }
}

// ------------------------------------------------------------------------------------------------
// VARIANT 3: the compact optimised permutation with every product by a round constant (matrix rows,
// sparse-row coefficients, lane updates) done through digit tables (const_dot): 64 N + 24 wide
// multiplies for an N-term row instead of 64 N + 64, and 88 instead of 128 for a lane update.  Only
// the S-box keeps true Montgomery products.  Same results, bit for bit.
// ------------------------------------------------------------------------------------------------
// out = sum_j st[j] * c_j with the table of c_0 at off_first and those of c_1.. at off_rest + 32 (j-1)
template <class F, int T>
PSD_DEV void row_cdot(uint32_t* out, const uint32_t (*st)[8], uint32_t off_first, uint32_t off_rest) {
    const_dot<F, T>(out, st, [=](int j, int dgt, uint32_t* t) {
        cload8(t, (j == 0 ? off_first : off_rest + (j - 1) * 32) + dgt * 8);
    });
}

// LANES_ONLY: only the lane updates go through digit tables (for widths whose full table set does not
// fit the constant pool, e.g. t = 5); rows stay lazily reduced Montgomery dot products as in variant 2.
template <class F, int T, int ALPHA, bool LANES_ONLY = false>
PSD_DEV void permute_opt_ctab(uint32_t (&st)[T][8], const PermDesc& d) {
    constexpr bool kLazySbox = !LANES_ONLY && F::kBits == 254 && (ALPHA == 5 || ALPHA == 17);
    const uint32_t first_partial = d.half_full, first_post = d.half_full + d.partial;
#if defined(__CUDACC__)
#pragma unroll 1
#endif
    for (uint32_t r = 0; r < d.total; r++) {
        const bool full = (r < first_partial || r >= first_post);
        const uint32_t sp = d.sparse_off + (r - first_partial) * (2 * T) * 8;                  // k lives here
        // w / v tables (LANES_ONLY: v tables only)
        const uint32_t spt = d.ct_sparse_off + (r - first_partial) * ((LANES_ONLY ? 1 : 2) * (T - 1)) * 32;
        const uint32_t ark = full ? ((r == first_post && d.partial) ? d.post_ark_off : d.ark_off + r * T * 8) : sp;
        const uint32_t mat = LANES_ONLY ? ((r + 1 == first_partial && d.partial) ? d.pre_mds_off : d.mds_off)
                                        : ((r + 1 == first_partial && d.partial) ? d.ct_pre_off : d.ct_mds_off);
        const uint32_t n_iter = full ? T : 1;
        // phase A: S-boxes
#if defined(__CUDACC__)
#pragma unroll 1
#endif
        for (uint32_t i = 0; i < n_iter; i++) {
            uint32_t c[8];
            cload8(c, ark + i * 8);
            fp_add<F>(st[0], st[0], c);
            // every consumer of an S-box output below is a const_dot unless LANES_ONLY
            sbox_compact<F, ALPHA, kLazySbox>(st[0], d);
            if (full) {
#pragma unroll
                for (int k = 0; k < 8; k++) {
                    const uint32_t t0 = st[0][k];
#pragma unroll
                    for (int l = 0; l + 1 < T; l++) st[l][k] = st[l + 1][k];
                    st[T - 1][k] = t0;
                }
            }
        }
        // phase B: rows
        uint32_t nw[T][8];
#pragma unroll
        for (int l = 0; l < T; l++)
#pragma unroll
            for (int k = 0; k < 8; k++) nw[l][k] = 0;
#if defined(__CUDACC__)
#pragma unroll 1
#endif
        for (uint32_t i = 0; i < n_iter; i++) {
            uint32_t t[8];
            if (LANES_ONLY) {
                row_dot<F, T>(t, st, full ? mat + i * T * 8 : sp + 8);
            } else {
                const uint32_t off_first = full ? mat + i * T * 32 : d.ct_m00_off;
                const uint32_t off_rest = full ? off_first + 32 : spt;
                row_cdot<F, T>(t, st, off_first, off_rest);
            }
#pragma unroll
            for (int k = 0; k < 8; k++) {
#pragma unroll
                for (int l = 0; l + 1 < T; l++) nw[l][k] = nw[l + 1][k];
                nw[T - 1][k] = t[k];
            }
        }
        if (full) {
#pragma unroll
            for (int l = 0; l < T; l++)
#pragma unroll
                for (int k = 0; k < 8; k++) st[l][k] = nw[l][k];
        } else {
            // phase C: lane updates s_j += v_j * s0
#if defined(__CUDACC__)
#pragma unroll 1
#endif
            for (uint32_t j = 0; j + 1 < T; j++) {
                uint32_t prod[8];
                const uint32_t voff = spt + ((LANES_ONLY ? 0 : T - 1) + j) * 32;
                const_dot<F, 1>(prod, st, [=](int, int dgt, uint32_t* t) { cload8(t, voff + dgt * 8); });
                fp_add<F>(st[1], st[1], prod);
#pragma unroll
                for (int k = 0; k < 8; k++) {
                    const uint32_t t1 = st[1][k];
#pragma unroll
                    for (int l = 1; l + 1 < T; l++) st[l][k] = st[l + 1][k];
                    st[T - 1][k] = t1;
                }
            }
#pragma unroll
            for (int k = 0; k < 8; k++) st[0][k] = nw[T - 1][k];
        }
    }
}

// Compact form for wide states (t >= 6): same operation order, loops kept rolled and the state
// indexed dynamically (it then lives in local memory).  These widths are not benchmark
// configurations; the point is bounded code size and compile time.
template <class F, int T, int ALPHA>
PSD_DEV void permute_rolled(uint32_t (&st)[T][8], const PermDesc& d) {
#if defined(__CUDACC__)
#pragma unroll 1
#endif
    for (uint32_t r = 0; r < d.total; r++) {
        const bool full = (r < d.half_full) || (r >= d.half_full + d.partial);
#if defined(__CUDACC__)
#pragma unroll 1
#endif
        for (int i = 0; i < T; i++) {
            uint32_t c[8];
            cload8(c, d.ark_off + (r * T + i) * 8);
            fp_add<F>(st[i], st[i], c);
        }
        const int n_sbox = full ? T : 1;
#if defined(__CUDACC__)
#pragma unroll 1
#endif
        for (int i = 0; i < n_sbox; i++) sbox<F, ALPHA>(st[i], d);
        uint32_t nw[T][8];
#if defined(__CUDACC__)
#pragma unroll 1
#endif
        for (int i = 0; i < T; i++) {
            uint32_t acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
#if defined(__CUDACC__)
#pragma unroll 1
#endif
            for (int j = 0; j < T; j++) {
                uint32_t m[8], prod[8];
                cload8(m, d.mds_off + (i * T + j) * 8);
                mont_mul<F>(prod, st[j], m);
                fp_add<F>(acc, acc, prod);
            }
#pragma unroll
            for (int k = 0; k < 8; k++) nw[i][k] = acc[k];
        }
#if defined(__CUDACC__)
#pragma unroll 1
#endif
        for (int i = 0; i < T; i++)
#pragma unroll
            for (int k = 0; k < 8; k++) st[i][k] = nw[i][k];
    }
}

// canonical <-> Montgomery (ark-ff from_bigint / into_bigint)
template <class F>
PSD_DEV void to_mont(uint32_t* x, const PermDesc& d) {
    uint32_t r2[8];
    cload8(r2, d.r2_off);
    mont_mul<F>(x, x, r2);
}
template <class F>
PSD_DEV void from_mont(uint32_t* x) {
    const uint32_t one[8] = {1, 0, 0, 0, 0, 0, 0, 0};
    mont_mul<F>(x, x, one);
}

// x < p ?  (canonical inputs must be reduced: ark-ff from_bigint -> None otherwise)
template <class F>
PSD_DEV bool is_canonical(const uint32_t* x) {
    const uint32_t p[8] = {F::P0, F::P1, F::P2, F::P3, F::P4, F::P5, F::P6, F::P7};
    bool lt = false, decided = false;
#pragma unroll
    for (int k = 7; k >= 0; k--) {
        if (!decided && x[k] != p[k]) {
            lt = x[k] < p[k];
            decided = true;
        }
    }
    return lt;
}

}  // namespace psd
