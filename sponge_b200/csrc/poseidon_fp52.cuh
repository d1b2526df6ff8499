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
// Bounds (5-limb values): lanes < 2^258 stay < 2^206 + 6p < 2^258; every limb < 2^52 after normalising.
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

template <class F, int T, int ALPHA>
PSD_DEV void permute_fp52(uint32_t (&st)[T][8], const PermDesc& d) {
    static_assert(F::kBits == 254 && (ALPHA == 5 || ALPHA == 17), "variant 5: BN254 Fr, alpha 5 or 17");
    uint64_t S[T][5];
    // round-0 constants on the integer side, then into limb form
#pragma unroll
    for (int j = 0; j < T; j++) {
        uint32_t c[8];
        cload8(c, d.ark_off + j * 8);
        fp_add<F>(st[j], st[j], c);
        f52_unpack(S[j], st[j]);
    }
    const uint32_t first_partial = d.half_full, first_post = d.half_full + d.partial;
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
    uint32_t k_off = d.fp_k_off, part_off = d.fp_part_off;
#if defined(__CUDACC__)
#pragma unroll 1
#endif
    for (uint32_t r = 0; r < d.total; r++) {
        const bool full = (r < first_partial || r >= first_post);
        const bool all_rows = full || (r + 1 == first_post);      // the last partial round runs T rows
        const uint32_t n_sbox = full ? T : 1, n_rows = all_rows ? T : 1;
        const uint32_t mat = full ? ((r + 1 == first_partial && d.partial) ? d.fp_pre_off : d.fp_mat_off)
                                  : d.fp_last_off;
        // phase A: S-boxes on lane 0, lanes rotated through it in the full rounds
#if defined(__CUDACC__)
#pragma unroll 1
#endif
        for (uint32_t i = 0; i < n_sbox; i++) {
            uint64_t plain[5];
#pragma unroll
            for (int k = 0; k < 5; k++) plain[k] = S[0][k] & kMask52;
            uint32_t x[8];
            f52_pack(x, plain);
            sbox_compact<F, ALPHA, true>(x, d);
            f52_unpack(S[0], x);
            if (full) {
#pragma unroll
                for (int k = 0; k < 5; k++) {
                    const uint64_t t0 = S[0][k];
#pragma unroll
                    for (int l = 0; l + 1 < T; l++) S[l][k] = S[l + 1][k];
                    S[T - 1][k] = t0;
                }
            }
        }
        // phase B: rows
        uint64_t nw[T][5];
#pragma unroll
        for (int l = 0; l < T; l++)
#pragma unroll
            for (int k = 0; k < 5; k++) nw[l][k] = 0;
#if defined(__CUDACC__)
#pragma unroll 1
#endif
        for (uint32_t i = 0; i < n_rows; i++) {
            uint64_t t[5];
            const uint32_t off_first = all_rows ? mat + i * T * kF52TabWords : d.fp_mat_off;
            const uint32_t off_rest = all_rows ? off_first + kF52TabWords : part_off;
            f52_row<T>(t, S, off_first, off_rest, k_off, consts, c104);
            k_off += kF52KWords;
#pragma unroll
            for (int k = 0; k < 5; k++) {
#pragma unroll
                for (int l = 0; l + 1 < T; l++) nw[l][k] = nw[l + 1][k];
                nw[T - 1][k] = t[k] | kMagicLo;
            }
        }
        if (all_rows) {
#pragma unroll
            for (int l = 0; l < T; l++)
#pragma unroll
                for (int k = 0; k < 5; k++) S[l][k] = nw[l][k];
        } else {
            // phase C: lane updates with y = S[0]
            double y[5];
#pragma unroll
            for (int k = 0; k < 5; k++) y[k] = f52_value(S[0][k]);
#if defined(__CUDACC__)
#pragma unroll 1
#endif
            for (uint32_t j = 0; j + 1 < T; j++) {
                f52_lane_update(S[1], y, part_off + (T - 1 + j) * kF52TabWords, consts, c104);
#pragma unroll
                for (int k = 0; k < 5; k++) {
                    const uint64_t t1 = S[1][k];
#pragma unroll
                    for (int l = 1; l + 1 < T; l++) S[l][k] = S[l + 1][k];
                    S[T - 1][k] = t1;
                }
            }
            part_off += 2 * (T - 1) * kF52TabWords;
#pragma unroll
            for (int k = 0; k < 5; k++) S[0][k] = nw[T - 1][k];
        }
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
