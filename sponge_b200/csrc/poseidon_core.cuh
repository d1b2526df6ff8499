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

#if defined(__CUDACC__)
__constant__ uint32_t c_pool[kConstPoolWords];
#else
static uint32_t c_pool[kConstPoolWords];
#endif

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
};

template <class F, int ALPHA>
PSD_DEV void sbox(uint32_t* x, const PermDesc& d) {
    if (ALPHA == 5) {
        fp_pow5<F>(x, x);
    } else if (ALPHA == 17) {
        uint32_t y[8];
        mont_mul<F>(y, x, x);
        mont_mul<F>(y, y, y);
        mont_mul<F>(y, y, y);
        mont_mul<F>(y, y, y);
        mont_mul<F>(x, y, x);
    } else {
        const uint64_t alpha = ((uint64_t)d.alpha_hi << 32) | d.alpha_lo;
        uint32_t one[8];
#pragma unroll
        for (int k = 0; k < 8; k++) one[k] = c_pool[d.one_off + k];
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
#pragma unroll
        for (int k = 0; k < 8; k++) c[k] = c_pool[d.ark_off + (r * T + i) * 8 + k];
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
#pragma unroll
            for (int k = 0; k < 8; k++) m[k] = c_pool[d.mds_off + (i * T + j) * 8 + k];
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
#pragma unroll
            for (int k = 0; k < 8; k++) c[k] = c_pool[d.ark_off + (r * T + i) * 8 + k];
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
#pragma unroll
                for (int k = 0; k < 8; k++) m[k] = c_pool[d.mds_off + (i * T + j) * 8 + k];
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
#pragma unroll
    for (int k = 0; k < 8; k++) r2[k] = c_pool[d.r2_off + k];
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
