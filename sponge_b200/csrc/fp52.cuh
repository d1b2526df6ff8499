// Products of field elements with ROUND CONSTANTS on the FP64 pipe (sm_100a: 64 DFMA/clk/SM, idle in
// the integer kernels), in 5 x 52-bit limbs held in doubles.
//
// Why: the integer multiplier (IMAD.WIDE, 16 results/clk/SM) is the roofline of variants 0-4 and is
// >90 % busy.  The linear layer of Poseidon (MDS rows, sparse rows, lane updates; poseidon/mod.rs:82-93
// after the regrouping of opt_constants.cpp) only multiplies the state by constants, which needs no
// general Montgomery product: for a constant c the host stores T_i = c * 2^(52 i) * 2^(52 r) mod p and
//     x * c  ==  ( sum_i X_i * T_i  +  (r Montgomery rows of 52 bits) ) / 2^(52 r)      (mod p)
// with X_i the 52-bit limbs of x.  Every X_i * T_i[k] is a 52x52-bit product done EXACTLY with two
// fused multiply-adds in round-toward-zero mode:
//     hi = fma_rz(a, b, 2^104)               -> 2^104 + floor(ab / 2^52) * 2^52     (mantissa field == H)
//     lo = fma_rz(a, b, (2^104 + 2^52) - hi) -> 2^52 + (ab mod 2^52)               (mantissa field == L)
// so the raw IEEE bit patterns of hi and lo are "magic + H" and "magic + L" and column sums are
// accumulated with plain 64-bit integer adds of those bit patterns (the magic exponents add up to a
// compile-time constant that the host folds into the additive round constant).  Per 52x52 product:
// 3 FP64-pipe instructions (2 DFMA + 1 DADD) and 2 ALU instructions (two 3-input 64-bit adds cover two
// products), no integer multiplies.  The S-box keeps the integer Montgomery code of fp256.cuh, so both
// multiplier pipes work at the same time.
//
// The same header compiles for the host (PSD_HOST_EMU) with fma() under FE_TOWARDZERO; tests/emu sets
// the rounding mode around every call.
#pragma once
#include "poseidon_core.cuh"

#if defined(PSD_HOST_EMU)
#include <math.h>
#include <string.h>
#endif

namespace psd {

constexpr uint64_t kMask52 = (1ull << 52) - 1;
constexpr uint64_t kMagicLo = 0x4330000000000000ull;   // bits of 2^52
constexpr uint64_t kMagicHi = 0x4670000000000000ull;   // bits of 2^104

#if defined(PSD_HOST_EMU)
static inline double f52_from_bits(uint64_t b) { double d; memcpy(&d, &b, 8); return d; }
static inline uint64_t f52_bits(double d) { uint64_t b; memcpy(&b, &d, 8); return b; }
static inline double f52_fma(double a, double b, double c) { return fma(a, b, c); }
static inline double f52_sub(double a, double b) { volatile double r = a - b; return r; }
#else
PSD_DEV double f52_from_bits(uint64_t b) { return __longlong_as_double((long long)b); }
PSD_DEV uint64_t f52_bits(double d) { return (uint64_t)__double_as_longlong(d); }
PSD_DEV double f52_fma(double a, double b, double c) { return __fma_rz(a, b, c); }
PSD_DEV double f52_sub(double a, double b) { return __dsub_rz(a, b); }
#endif

// The pool as doubles / int64 bit patterns; offsets count 8-byte units.  Tables are addressed as
// pointer + compile-time index so that every read is one LDCU.64 at base + immediate.
PSD_DEV const double* f52_pool_d(uint32_t off) { return reinterpret_cast<const double*>(c_pool) + off; }
PSD_DEV const uint64_t* f52_pool_u(uint32_t off) { return reinterpret_cast<const uint64_t*>(c_pool) + off; }

// "magic" form of a 52-bit limb X: the bit pattern of the double 2^52 + X.  The state lives in this
// form between rounds: as an integer it is X + kMagicLo (ready for the column sums), as a double it
// is one exact subtraction away from X.
PSD_DEV double f52_value(uint64_t magic) { return f52_sub(f52_from_bits(magic), 4503599627370496.0); }

// (hi, lo) bit patterns of the exact product a*b, a, b integers in [0, 2^52).  c104 = 2^104 is passed
// in so that it sits in a vector register: DFMA takes ONE uniform-register or immediate operand, and
// that slot is needed for the table constant b.
PSD_DEV void f52_prod(uint64_t& hi_bits, uint64_t& lo_bits, double a, double b, double c104) {
    const double hi = f52_fma(a, b, c104);                                          // 2^104 + H 2^52
    const double sub = f52_sub(20282409603651674927546878656512.0, hi);             // 2^104 + 2^52 - hi
    const double lo = f52_fma(a, b, sub);
    hi_bits = f52_bits(hi);
    lo_bits = f52_bits(lo);
}

// c[0..5] += a * tab[0..4]  (tab = five 52-bit limbs of one table entry, little endian)
PSD_DEV void f52_mac5(uint64_t* c, double a, const double* tab, double c104) {
    uint64_t hi[5], lo[5];
#pragma unroll
    for (int k = 0; k < 5; k++) f52_prod(hi[k], lo[k], a, tab[k], c104);
    c[0] += lo[0];
#pragma unroll
    for (int k = 1; k < 5; k++) c[k] += lo[k] + hi[k - 1];
    c[5] += hi[4];
}

// Layout of the FP constants of one parameter set (8-byte units into the pool), see fp52_tables.cpp.
//   consts: P52[0..4] (doubles), NP52 (uint64: -p^-1 mod 2^52), SHIFT (uint64, see f52_mont_row), 2^104 (double)
constexpr uint32_t kF52ConstP = 0, kF52ConstNp = 5, kF52ConstShift = 6, kF52ConstC104 = 7, kF52ConstWords = 8;

// One 52-bit Montgomery row on the column sums: c <- (c + m p) / 2^52 with m = -c/p mod 2^52.
// The magic offsets of the 5 products are already inside the columns (folded by the host); `shift_c5`
// is the value the vacated top column starts from (minus the offsets it will receive).
PSD_DEV void f52_mont_row(uint64_t* c, const double* consts, uint64_t shift_c5, double c104) {
    const uint64_t q = c[0] & kMask52;
    const uint64_t np = reinterpret_cast<const uint64_t*>(consts)[kF52ConstNp];
    const uint64_t m = (q * np) & kMask52;
    const double md = f52_value(m | kMagicLo);
    f52_mac5(c, md, consts + kF52ConstP, c104);
    // the low 52 bits of c[0] are now zero
    c[1] += c[0] >> 52;
#pragma unroll
    for (int k = 0; k < 5; k++) c[k] = c[k + 1];
    c[5] = shift_c5;
}

// carries of columns 0..3 pushed up; afterwards every c[k] < 2^52 (c[4] by the caller's bound)
PSD_DEV void f52_normalise(uint64_t* c) {
#pragma unroll
    for (int k = 0; k < 4; k++) {
        c[k + 1] += c[k] >> 52;
        c[k] &= kMask52;
    }
}

// 8 x 32-bit words (value < 2^256) -> five magic limbs
PSD_DEV void f52_unpack(uint64_t* m, const uint32_t* x) {
    const uint64_t w01 = (uint64_t)x[0] | ((uint64_t)x[1] << 32);
    const uint64_t w12 = (uint64_t)x[1] | ((uint64_t)x[2] << 32);
    const uint64_t w34 = (uint64_t)x[3] | ((uint64_t)x[4] << 32);
    const uint64_t w45 = (uint64_t)x[4] | ((uint64_t)x[5] << 32);
    const uint64_t w67 = (uint64_t)x[6] | ((uint64_t)x[7] << 32);
    m[0] = (w01 & kMask52) | kMagicLo;                                              // bits   0..51
    m[1] = (((w12 >> 20) | ((uint64_t)x[3] << 44)) & kMask52) | kMagicLo;           // bits  52..103
    m[2] = ((w34 >> 8) & kMask52) | kMagicLo;                                       // bits 104..155
    m[3] = (((w45 >> 28) | ((uint64_t)x[6] << 36)) & kMask52) | kMagicLo;           // bits 156..207
    m[4] = (w67 >> 16) | kMagicLo;                                                  // bits 208..255
}

// five normalised limbs (plain integers < 2^52, value < 2^256) -> 8 x 32-bit words
PSD_DEV void f52_pack(uint32_t* x, const uint64_t* c) {
    x[0] = (uint32_t)c[0];
    x[1] = (uint32_t)(c[0] >> 32) | (uint32_t)(c[1] << 20);
    x[2] = (uint32_t)(c[1] >> 12);
    x[3] = (uint32_t)(c[1] >> 44) | (uint32_t)(c[2] << 8);
    x[4] = (uint32_t)(c[2] >> 24) | (uint32_t)(c[3] << 28);
    x[5] = (uint32_t)(c[3] >> 4);
    x[6] = (uint32_t)(c[3] >> 36) | (uint32_t)(c[4] << 16);
    x[7] = (uint32_t)(c[4] >> 16);
}

}  // namespace psd
