// VARIANT 5: the optimised Poseidon permutation with the S-box on the integer multiplier and the whole
// linear layer on the FP64 pipe (fp52.cuh).  Same function, bit for bit, as poseidon/mod.rs:95-118.
//
// State between rounds: T lanes of five "magic" limbs (bit pattern of 2^52 + X_i, see fp52.cuh).
//   phase A  S-box lanes: pack to 8 x 32-bit words, x^alpha with the integer Montgomery code (lazy form:
//            inputs < 1.01 p, outputs < 1.26 p, any 256-bit value is a valid multiplicand below), unpack.
//            The additive round constant is NOT added here: it was folded into the row that produced x.
//   phase B  rows: out = (K + sum_j sum_i X_{j,i} * T_{j,i} + two Montgomery rows) / 2^104  -> < p + 2^207.
//   phase C  partial rounds: lanes s_j <- (s_j + sum_i Y_i * TV_{j,i} + one Montgomery row) / 2^52.
// Lane scaling.  Phase C divides by 2^52 instead of reducing fully, so in the partial rounds lanes
// 1..T-1 hold s_j * sigma_i with sigma_i = 2^(-52 i) after i lane updates; the host scales the tables
// of partial round i accordingly (w_j / sigma_i, v_j * sigma_i), and the last partial round is run as
// T full rows whose matrix also undoes the scale, so the full rounds never see it.
// Bounds (5-limb values): lanes < 2^258 stay < 2^206 + 6p < 2^258 (6p = 1.13 * 2^256 for BN254, 2.72 * 2^256 for
// BLS12-381); every limb < 2^52 after normalising.  For BLS12-381 the S-box input is brought below p first
// (one conditional subtraction) and the S-box keeps its reductions.
#pragma once
#include "fp52.cuh"

namespace psd {

constexpr uint32_t kF52TabWords = 25;   // one table: 5 entries x 5 limbs (doubles)
constexpr uint32_t kF52KWords = 6;      // one additive constant: 6 column initial values

// Magic offsets that reach each column of a lane update (1 lane + 5 products + 1 Montgomery row):
// column 0 receives 7 "lo" patterns, columns 1..4 7 lo + 6 hi, column 5 6 hi.
constexpr uint64_t kF52LaneC0 = 0 - 7 * kMagicLo;
constexpr uint64_t kF52LaneCk = 0 - (7 * kMagicLo + 6 * kMagicHi);
constexpr uint64_t kF52LaneC5 = 0 - 6 * kMagicHi;

// out (five normalised plain limbs) = row over the T lanes of S.  Table of lane 0 at off_first, of
// lane j >= 1 at off_rest + 25 (j - 1); additive constant (with every magic offset folded in) at k_off.
template <int T>
PSD_DEV void f52_row(uint64_t* out, const uint64_t (*S)[5], uint32_t off_first, uint32_t off_rest, uint32_t k_off,
                     const double* consts, double c104) {
    uint64_t c[6];
    const uint64_t* kk = f52_pool_u(k_off);
#pragma unroll
    for (int k = 0; k < 6; k++) c[k] = kk[k];
#pragma unroll
    for (int j = 0; j < T; j++) {
        const double* tab = f52_pool_d((j == 0) ? off_first : off_rest + (j - 1) * kF52TabWords);
#pragma unroll
        for (int i = 0; i < 5; i++) f52_mac5(c, f52_value(S[j][i]), tab + i * 5, c104);
    }
    const uint64_t shift = reinterpret_cast<const uint64_t*>(consts)[kF52ConstShift];
    f52_mont_row(c, consts, shift, c104);
    f52_mont_row(c, consts, shift, c104);
    f52_normalise(c);
#pragma unroll
    for (int k = 0; k < 5; k++) out[k] = c[k];
}

// lane (magic limbs) <- (lane + y * v) / 2^52, y given as doubles, v by its table
PSD_DEV void f52_lane_update(uint64_t* lane, const double* y, uint32_t tab_off, const double* consts, double c104) {
    uint64_t c[6];
    c[0] = lane[0] + kF52LaneC0;
#pragma unroll
    for (int k = 1; k < 5; k++) c[k] = lane[k] + kF52LaneCk;
    c[5] = kF52LaneC5;
    const double* tab = f52_pool_d(tab_off);
#pragma unroll
    for (int i = 0; i < 5; i++) f52_mac5(c, y[i], tab + i * 5, c104);
    f52_mont_row(c, consts, 0, c104);
    f52_normalise(c);
#pragma unroll
    for (int k = 0; k < 5; k++) lane[k] = c[k] | kMagicLo;
}

// c = K + sum over lanes j >= 1 of lane_j * TW_j: the part of a partial round's row that does not
// depend on the S-box output of that round.
template <int T>
PSD_DEV void f52_row_lanes(uint64_t* c, const uint64_t (*S)[5], uint32_t tw_off, uint32_t k_off, double c104) {
    const uint64_t* kk = f52_pool_u(k_off);
#pragma unroll
    for (int k = 0; k < 6; k++) c[k] = kk[k];
#pragma unroll
    for (int j = 1; j < T; j++) {
        const double* tab = f52_pool_d(tw_off + (j - 1) * kF52TabWords);
#pragma unroll
        for (int i = 0; i < 5; i++) f52_mac5(c, f52_value(S[j][i]), tab + i * 5, c104);
    }
}

// Which fields may skip the conditional subtractions inside x^alpha (see sbox_compact<.., LAZY>): BN254
// (5.29 p > 2^256).  BLS12-381 (2p > 2^255, and its squaring folds a to min(a, p-a)) needs canonical
// S-box inputs and produces canonical outputs.
template <class F>
struct F52Lazy {
    static constexpr bool value = (F::kBits == 254);
};

// x (five plain limbs, value < p + 2^207) -> S-box input words
template <class F>
PSD_DEV void f52_pack_sbox_input(uint32_t* x, const uint64_t* plain) {
    if (F52Lazy<F>::value) {
        f52_pack(x, plain);
    } else {
        uint32_t w[8];
        f52_pack(w, plain);
        cond_sub_p<F>(x, w);
    }
}

// x^alpha, straight-line (no loop, so that it shares a basic block with the FP64 work beside it)
template <class F, int ALPHA>
PSD_DEV void f52_sbox_straight(uint32_t* x) {
    constexpr bool kReduce = !F52Lazy<F>::value;
    uint32_t y[8];
    fp_sqr<F, kReduce>(y, x);
#pragma unroll
    for (int q = 1; q < (ALPHA == 5 ? 2 : 4); q++) fp_sqr<F, kReduce>(y, y);
    mont_mul<F, kReduce>(x, y, x);
}

// One full round on the limb state: T S-boxes (lane 0, lanes rotated through it), then T rows.
template <class F, int T, int ALPHA>
PSD_DEV void f52_full_round(uint64_t (&S)[T][5], uint32_t mat, uint32_t first_row, uint32_t k_off, const double* consts,
                            double c104, const PermDesc& d, bool sboxes) {
    if (sboxes) {
#if defined(__CUDACC__)
#pragma unroll 1
#endif
        for (uint32_t i = 0; i < T; i++) {
            uint64_t plain[5];
#pragma unroll
            for (int k = 0; k < 5; k++) plain[k] = S[0][k] & kMask52;
            uint32_t x[8];
            f52_pack_sbox_input<F>(x, plain);
            sbox_compact<F, ALPHA, F52Lazy<F>::value>(x, d);
            f52_unpack(S[0], x);
#pragma unroll
            for (int k = 0; k < 5; k++) {
                const uint64_t t0 = S[0][k];
#pragma unroll
                for (int l = 0; l + 1 < T; l++) S[l][k] = S[l + 1][k];
                S[T - 1][k] = t0;
            }
        }
    }
    // rows first_row..T-1 (rows below first_row keep their lanes: nw starts as a copy shifted in below)
    uint64_t nw[T][5];
#pragma unroll
    for (int l = 0; l < T; l++)
#pragma unroll
        for (int k = 0; k < 5; k++) nw[l][k] = 0;
#if defined(__CUDACC__)
#pragma unroll 1
#endif
    for (uint32_t i = first_row; i < T; i++) {
        uint64_t t[5];
        const uint32_t off_first = mat + i * T * kF52TabWords;
        f52_row<T>(t, S, off_first, off_first + kF52TabWords, k_off + (i - first_row) * kF52KWords, consts, c104);
#pragma unroll
        for (int k = 0; k < 5; k++) {
#pragma unroll
            for (int l = 0; l + 1 < T; l++) nw[l][k] = nw[l + 1][k];
            nw[T - 1][k] = t[k] | kMagicLo;
        }
    }
    // after T - first_row shifts, row i sits in nw[i]; rows below first_row are the caller's business
#pragma unroll
    for (int l = 0; l < T; l++)
#pragma unroll
        for (int k = 0; k < 5; k++)
            if (l >= 1 || first_row == 0) S[l][k] = nw[l][k];
}

template <class F, int T, int ALPHA>
PSD_DEV void permute_fp52(uint32_t (&st)[T][8], const PermDesc& d) {
    static_assert(ALPHA == 5 || ALPHA == 17, "variant 5: alpha 5 or 17");
    uint64_t S[T][5];
    // round-0 constants on the integer side, then into limb form
#pragma unroll
    for (int j = 0; j < T; j++) {
        uint32_t c[8];
        cload8(c, d.ark_off + j * 8);
        fp_add<F>(st[j], st[j], c);
        f52_unpack(S[j], st[j]);
    }
    const uint32_t half = d.half_full, rp = d.partial;
    const double* consts = f52_pool_d(d.fp_consts_off);
    // 2^104 must sit in a VECTOR register (see f52_prod); a value the compiler can prove warp-uniform
    // would be kept in a uniform register or folded into an immediate, and either way each product
    // would pay two extra moves.  The active mask of a running thread is never zero, which the
    // compiler cannot know.
#if defined(__CUDACC__)
    const double c104 = __hiloint2double((int)(0x46700000u + (__activemask() == 0u ? 1u : 0u)), 0);
#else
    const double c104 = consts[kF52ConstC104];
#endif
    const uint64_t shift = reinterpret_cast<const uint64_t*>(consts)[kF52ConstShift];
    uint32_t k_off = d.fp_k_off;

    // full rounds before the partial rounds
#if defined(__CUDACC__)
#pragma unroll 1
#endif
    for (uint32_t r = 0; r < half; r++) {
        f52_full_round<F, T, ALPHA>(S, (r + 1 == half && rp) ? d.fp_pre_off : d.fp_mat_off, 0, k_off, consts, c104, d, true);
        k_off += T * kF52KWords;
    }

    if (rp) {
        // Partial rounds.  Only lane 0's chain  S-box -> row 0 -> S-box  is serial; the lane updates of
        // round i and the lanes' share of row 0 of round i+1 depend on nothing the S-box of round i+1
        // produces, so they are written next to it; the S-box output's share of the row and the two
        // Montgomery rows follow once the S-box is done.
        uint32_t part_off = d.fp_part_off;
        uint32_t x[8];
        {
            uint64_t plain[5];
#pragma unroll
            for (int k = 0; k < 5; k++) plain[k] = S[0][k] & kMask52;
            f52_pack_sbox_input<F>(x, plain);
        }
        uint64_t c[6];
        f52_sbox_straight<F, ALPHA>(x);
        f52_row_lanes<T>(c, S, part_off, k_off, c104);
#if defined(__CUDACC__)
#pragma unroll 1
#endif
        for (uint32_t i = 0;; i++) {
            // finish row 0 of partial round i: the S-box output's share, two Montgomery rows
            f52_unpack(S[0], x);
            double y[5];
#pragma unroll
            for (int k = 0; k < 5; k++) y[k] = f52_value(S[0][k]);
            {
                const double* tab = f52_pool_d(d.fp_mat_off);
#pragma unroll
                for (int l = 0; l < 5; l++) f52_mac5(c, y[l], tab + l * 5, c104);
            }
            f52_mont_row(c, consts, shift, c104);
            f52_mont_row(c, consts, shift, c104);
            f52_normalise(c);
            k_off += kF52KWords;
            if (i + 1 == rp) break;
            f52_pack_sbox_input<F>(x, c);
            // S-box of round i+1  ||  lane updates of round i, lane part of row 0 of round i+1: one
            // straight-line block.  Measured on B200 (c2): 163.1 M perm/s, against 162.2 M for rolled
            // phases in program order and 151-157 M for a rolled lane loop next to a rolled S-box, even
            // though this body (1.5 K instructions) misses the instruction cache (ncu: no_instruction
            // is 18 % of the stall samples).
            f52_sbox_straight<F, ALPHA>(x);
#pragma unroll
            for (int j = 1; j < T; j++)
                f52_lane_update(S[j], y, part_off + (T - 1 + j - 1) * kF52TabWords, consts, c104);
            part_off += 2 * (T - 1) * kF52TabWords;
            f52_row_lanes<T>(c, S, part_off, k_off, c104);
        }
        // last partial round: lanes 1..T-1 come from rows of the `last` matrix (lane update + rescale),
        // lane 0 is the row just finished
        uint64_t lane0[5];
#pragma unroll
        for (int k = 0; k < 5; k++) lane0[k] = c[k] | kMagicLo;
        f52_full_round<F, T, ALPHA>(S, d.fp_last_off, 1, k_off, consts, c104, d, false);
        k_off += (T - 1) * kF52KWords;
#pragma unroll
        for (int k = 0; k < 5; k++) S[0][k] = lane0[k];
    }

    // full rounds after the partial rounds
#if defined(__CUDACC__)
#pragma unroll 1
#endif
    for (uint32_t r = half + rp; r < d.total; r++) {
        f52_full_round<F, T, ALPHA>(S, d.fp_mat_off, 0, k_off, consts, c104, d, true);
        k_off += T * kF52KWords;
    }

    // back to canonical 8 x 32-bit residues (rows leave values < p + 2^207)
#pragma unroll
    for (int j = 0; j < T; j++) {
        uint64_t plain[5];
#pragma unroll
        for (int k = 0; k < 5; k++) plain[k] = S[j][k] & kMask52;
        uint32_t x[8];
        f52_pack(x, plain);
        cond_sub_p<F>(st[j], x);
    }
}

}  // namespace psd
