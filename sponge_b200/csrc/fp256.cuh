// 256-bit Montgomery arithmetic over BN254 Fr and BLS12-381 Fr in 8 x 32-bit limbs (sm_100a).
//
// Replaces, on the device, the `ark_ff::Fp<MontBackend<_,4>,4>` operations that the reference's
// Poseidon permutation bottoms out in (`add_assign`, `mul`, `pow`; call sites
// /root/reference/src/poseidon/mod.rs:67,72,78,85,87,88,128,143).  Values are Montgomery
// residues x*2^256 mod p; with R = 2^256 the 4x64-bit (ark-ff) and 8x32-bit (here) limb views are
// the same bytes.
//
// Multiplication is a CIOS Montgomery product whose 32x32->64 partial products are accumulated in
// two interleaved 64-bit-aligned accumulators ("lo-aligned" columns 0,2,4,6 and "hi-aligned" columns
// 1,3,5,7).  Each group of four products is one PTX carry chain (mad.lo.cc / madc.hi.cc), which
// ptxas fuses into IMAD.WIDE.U32 / IMAD.WIDE.U32.X with the carry in a predicate register: one
// integer-multiply-pipe instruction per 32x32 product and almost no separate carry arithmetic.
// Every chain is a single asm statement so the compiler can never separate a carry producer from
// its consumer.
//
// The same header compiles with plain g++ (PSD_HOST_EMU): each chain then runs a C model with an
// explicit carry, which lets tests/ check the limb plumbing on a machine without a GPU.
#pragma once
#include <stdint.h>

#if defined(__CUDACC__)
#define PSD_DEV __device__ __forceinline__
#else
#define PSD_DEV static inline
#ifndef PSD_HOST_EMU
#define PSD_HOST_EMU 1
#endif
#endif

namespace psd {

// ------------------------------------------------------------------------------------------------
// Field descriptors.  P = modulus limbs (LE), NP0 = -p^-1 mod 2^32.
// BLS12-381 Fr modulus: /root/reference/src/test.rs:6.  BN254 Fr: not in the reference
// (BASELINE.json configs 1-3); the standard alt_bn128 scalar field.
// ------------------------------------------------------------------------------------------------
struct Bn254Fr {
    static constexpr int kFieldId = 0;
    static constexpr int kBits = 254;
    static constexpr uint32_t P0 = 0xf0000001u, P1 = 0x43e1f593u, P2 = 0x79b97091u, P3 = 0x2833e848u,
                              P4 = 0x8181585du, P5 = 0xb85045b6u, P6 = 0xe131a029u, P7 = 0x30644e72u;
    static constexpr uint32_t NP0 = 0xefffffffu;
    static constexpr bool kModulusFromConstBank = false;
};
struct Bls12381Fr {
    static constexpr int kFieldId = 1;
    static constexpr int kBits = 255;
    static constexpr uint32_t P0 = 0x00000001u, P1 = 0xffffffffu, P2 = 0xfffe5bfeu, P3 = 0x53bda402u,
                              P4 = 0x09a1d805u, P5 = 0x3339d808u, P6 = 0x299d7d48u, P7 = 0x73eda753u;
    static constexpr uint32_t NP0 = 0xffffffffu;
    // With p0 = 1 and p1 = 2^32-1 as immediates ptxas rewrites m*p into IMAD.X / IMAD.HI.U32.X pairs
    // (6 multiplier-pipe cycles per product instead of 4); reading the limbs from the constant bank
    // keeps them opaque, so every product stays one IMAD.WIDE.U32.X with a uniform-register operand.
    static constexpr bool kModulusFromConstBank = true;
};

#if defined(__CUDACC__)
__constant__ uint32_t c_bls12_381_fr_p[9] = {0x00000001u, 0xffffffffu, 0xfffe5bfeu, 0x53bda402u,
                                             0x09a1d805u, 0x3339d808u, 0x299d7d48u, 0x73eda753u, 0xffffffffu};
#else
static const uint32_t c_bls12_381_fr_p[9] = {0x00000001u, 0xffffffffu, 0xfffe5bfeu, 0x53bda402u,
                                             0x09a1d805u, 0x3339d808u, 0x299d7d48u, 0x73eda753u, 0xffffffffu};
#endif

// -p^-1 mod 2^32 as the multiplier sees it
template <class F>
PSD_DEV uint32_t np0() {
    if (F::kModulusFromConstBank) return c_bls12_381_fr_p[8];
    return F::NP0;
}

// k-th modulus limb as the multiplier sees it (immediate or constant-bank operand)
template <class F, int K>
PSD_DEV uint32_t plimb() {
    if (F::kModulusFromConstBank) return c_bls12_381_fr_p[K];
    return K == 0 ? F::P0 : K == 1 ? F::P1 : K == 2 ? F::P2 : K == 3 ? F::P3 : K == 4 ? F::P4 : K == 5 ? F::P5
         : K == 6 ? F::P6 : F::P7;
}

// ------------------------------------------------------------------------------------------------
// Carry-chain primitives.  An accumulator is uint32_t[8] holding four (lo,hi) pairs.
// ------------------------------------------------------------------------------------------------

// d[2k],d[2k+1] = a_k * b   (four independent 32x32->64 products)
PSD_DEV void mul4(uint32_t* d, uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b) {
#if !defined(PSD_HOST_EMU)
    // mul.wide + unpack keeps each product a single IMAD.WIDE.U32 (mul.lo/mul.hi pairs are not fused)
    asm("{\n\t"
        ".reg .u64 t0, t1, t2, t3;\n\t"
        "mul.wide.u32 t0, %8, %12;\n\t"
        "mul.wide.u32 t1, %9, %12;\n\t"
        "mul.wide.u32 t2, %10, %12;\n\t"
        "mul.wide.u32 t3, %11, %12;\n\t"
        "mov.b64 {%0, %1}, t0;\n\t"
        "mov.b64 {%2, %3}, t1;\n\t"
        "mov.b64 {%4, %5}, t2;\n\t"
        "mov.b64 {%6, %7}, t3;\n\t"
        "}"
        : "=r"(d[0]), "=r"(d[1]), "=r"(d[2]), "=r"(d[3]), "=r"(d[4]), "=r"(d[5]), "=r"(d[6]), "=r"(d[7])
        : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b));
#else
    const uint32_t a[4] = {a0, a1, a2, a3};
    for (int k = 0; k < 4; k++) {
        uint64_t w = (uint64_t)a[k] * b;
        d[2 * k] = (uint32_t)w;
        d[2 * k + 1] = (uint32_t)(w >> 32);
    }
#endif
}

// d (256-bit, as 4 pairs) += {a0,a1,a2,a3} * b  laid out pairwise; the carry out of the top pair is
// added to `top`.
PSD_DEV void mad4(uint32_t* d, uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b,
                  uint32_t& top) {
#if !defined(PSD_HOST_EMU)
    asm("mad.lo.cc.u32 %0, %9, %13, %0;\n\t"
        "madc.hi.cc.u32 %1, %9, %13, %1;\n\t"
        "madc.lo.cc.u32 %2, %10, %13, %2;\n\t"
        "madc.hi.cc.u32 %3, %10, %13, %3;\n\t"
        "madc.lo.cc.u32 %4, %11, %13, %4;\n\t"
        "madc.hi.cc.u32 %5, %11, %13, %5;\n\t"
        "madc.lo.cc.u32 %6, %12, %13, %6;\n\t"
        "madc.hi.cc.u32 %7, %12, %13, %7;\n\t"
        "addc.u32 %8, %8, 0;"
        : "+r"(d[0]), "+r"(d[1]), "+r"(d[2]), "+r"(d[3]), "+r"(d[4]), "+r"(d[5]), "+r"(d[6]), "+r"(d[7]),
          "+r"(top)
        : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b));
#else
    const uint32_t a[4] = {a0, a1, a2, a3};
    uint64_t c = 0;
    for (int k = 0; k < 4; k++) {
        unsigned __int128 w = (unsigned __int128)((uint64_t)a[k] * b) +
                              (((uint64_t)d[2 * k + 1] << 32) | d[2 * k]) + c;
        d[2 * k] = (uint32_t)w;
        d[2 * k + 1] = (uint32_t)(w >> 32);
        c = (uint64_t)(w >> 64);
    }
    top += (uint32_t)c;
#endif
}

// Same, for chains whose carry out is provably zero (see the bound in mont_mul).
PSD_DEV void mad4_nc(uint32_t* d, uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b) {
#if !defined(PSD_HOST_EMU)
    asm("mad.lo.cc.u32 %0, %8, %12, %0;\n\t"
        "madc.hi.cc.u32 %1, %8, %12, %1;\n\t"
        "madc.lo.cc.u32 %2, %9, %12, %2;\n\t"
        "madc.hi.cc.u32 %3, %9, %12, %3;\n\t"
        "madc.lo.cc.u32 %4, %10, %12, %4;\n\t"
        "madc.hi.cc.u32 %5, %10, %12, %5;\n\t"
        "madc.lo.cc.u32 %6, %11, %12, %6;\n\t"
        "madc.hi.u32 %7, %11, %12, %7;"
        : "+r"(d[0]), "+r"(d[1]), "+r"(d[2]), "+r"(d[3]), "+r"(d[4]), "+r"(d[5]), "+r"(d[6]), "+r"(d[7])
        : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b));
#else
    uint32_t top = 0;
    mad4(d, a0, a1, a2, a3, b, top);
#endif
}

// The window slides down by one 32-bit column: `x` (whose x[0] is already zero) becomes the new
// hi-aligned accumulator and `y0` is column 0 of the new lo-aligned one.
//   y0 += x[1]                                  (carry feeds the chain below)
//   x[2k],x[2k+1] = a_k*b + (x[2k+2],x[2k+3])   k = 0..2
//   x[6],x[7]     = a_3*b + carry
PSD_DEV void mad4_slide(uint32_t* x, uint32_t& y0, uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3,
                        uint32_t b) {
#if !defined(PSD_HOST_EMU)
    asm("add.cc.u32 %8, %8, %1;\n\t"
        "madc.lo.cc.u32 %0, %9, %13, %2;\n\t"
        "madc.hi.cc.u32 %1, %9, %13, %3;\n\t"
        "madc.lo.cc.u32 %2, %10, %13, %4;\n\t"
        "madc.hi.cc.u32 %3, %10, %13, %5;\n\t"
        "madc.lo.cc.u32 %4, %11, %13, %6;\n\t"
        "madc.hi.cc.u32 %5, %11, %13, %7;\n\t"
        "madc.lo.cc.u32 %6, %12, %13, 0;\n\t"
        "madc.hi.u32 %7, %12, %13, 0;"
        : "+r"(x[0]), "+r"(x[1]), "+r"(x[2]), "+r"(x[3]), "+r"(x[4]), "+r"(x[5]), "+r"(x[6]), "+r"(x[7]),
          "+r"(y0)
        : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b));
#else
    const uint32_t a[4] = {a0, a1, a2, a3};
    uint64_t s = (uint64_t)y0 + x[1];
    y0 = (uint32_t)s;
    uint64_t c = s >> 32;
    for (int k = 0; k < 4; k++) {
        uint64_t add = (k < 3) ? ((((uint64_t)x[2 * k + 3]) << 32) | x[2 * k + 2]) : 0;
        unsigned __int128 w = (unsigned __int128)((uint64_t)a[k] * b) + add + c;
        x[2 * k] = (uint32_t)w;
        x[2 * k + 1] = (uint32_t)(w >> 32);
        c = (uint64_t)(w >> 64);
    }
#endif
}

// ---- partial chains for squaring: products for the first S pairs are omitted -------------------
// d pairs S..3 += {a_S..a_3} * b, carry out of the top pair added to `top`.  S == 4: nothing to do.
template <int S>
PSD_DEV void mad4_from(uint32_t* d, uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b,
                       uint32_t& top) {
    if (S <= 0) {
        mad4(d, a0, a1, a2, a3, b, top);
        return;
    }
    if (S >= 4) return;
#if !defined(PSD_HOST_EMU)
    if (S == 1) {
        asm("mad.lo.cc.u32 %0, %7, %10, %0;\n\t"
            "madc.hi.cc.u32 %1, %7, %10, %1;\n\t"
            "madc.lo.cc.u32 %2, %8, %10, %2;\n\t"
            "madc.hi.cc.u32 %3, %8, %10, %3;\n\t"
            "madc.lo.cc.u32 %4, %9, %10, %4;\n\t"
            "madc.hi.cc.u32 %5, %9, %10, %5;\n\t"
            "addc.u32 %6, %6, 0;"
            : "+r"(d[2]), "+r"(d[3]), "+r"(d[4]), "+r"(d[5]), "+r"(d[6]), "+r"(d[7]), "+r"(top)
            : "r"(a1), "r"(a2), "r"(a3), "r"(b));
    } else if (S == 2) {
        asm("mad.lo.cc.u32 %0, %5, %7, %0;\n\t"
            "madc.hi.cc.u32 %1, %5, %7, %1;\n\t"
            "madc.lo.cc.u32 %2, %6, %7, %2;\n\t"
            "madc.hi.cc.u32 %3, %6, %7, %3;\n\t"
            "addc.u32 %4, %4, 0;"
            : "+r"(d[4]), "+r"(d[5]), "+r"(d[6]), "+r"(d[7]), "+r"(top)
            : "r"(a2), "r"(a3), "r"(b));
    } else {
        asm("mad.lo.cc.u32 %0, %3, %4, %0;\n\t"
            "madc.hi.cc.u32 %1, %3, %4, %1;\n\t"
            "addc.u32 %2, %2, 0;"
            : "+r"(d[6]), "+r"(d[7]), "+r"(top)
            : "r"(a3), "r"(b));
    }
#else
    const uint32_t a[4] = {a0, a1, a2, a3};
    uint64_t c = 0;
    for (int k = S; k < 4; k++) {
        unsigned __int128 w = (unsigned __int128)((uint64_t)a[k] * b) +
                              (((uint64_t)d[2 * k + 1] << 32) | d[2 * k]) + c;
        d[2 * k] = (uint32_t)w;
        d[2 * k + 1] = (uint32_t)(w >> 32);
        c = (uint64_t)(w >> 64);
    }
    top += (uint32_t)c;
#endif
}

// mad4_slide with the products of the first S pairs omitted (those pairs are only shifted down and
// take the carry).  S in 0..3.
template <int S>
PSD_DEV void mad4_slide_from(uint32_t* x, uint32_t& y0, uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3,
                             uint32_t b) {
    if (S <= 0) {
        mad4_slide(x, y0, a0, a1, a2, a3, b);
        return;
    }
#if !defined(PSD_HOST_EMU)
    if (S == 1) {
        asm("add.cc.u32 %8, %8, %1;\n\t"
            "addc.cc.u32 %0, %2, 0;\n\t"
            "addc.cc.u32 %1, %3, 0;\n\t"
            "madc.lo.cc.u32 %2, %9, %12, %4;\n\t"
            "madc.hi.cc.u32 %3, %9, %12, %5;\n\t"
            "madc.lo.cc.u32 %4, %10, %12, %6;\n\t"
            "madc.hi.cc.u32 %5, %10, %12, %7;\n\t"
            "madc.lo.cc.u32 %6, %11, %12, 0;\n\t"
            "madc.hi.u32 %7, %11, %12, 0;"
            : "+r"(x[0]), "+r"(x[1]), "+r"(x[2]), "+r"(x[3]), "+r"(x[4]), "+r"(x[5]), "+r"(x[6]), "+r"(x[7]),
              "+r"(y0)
            : "r"(a1), "r"(a2), "r"(a3), "r"(b));
    } else if (S == 2) {
        asm("add.cc.u32 %8, %8, %1;\n\t"
            "addc.cc.u32 %0, %2, 0;\n\t"
            "addc.cc.u32 %1, %3, 0;\n\t"
            "addc.cc.u32 %2, %4, 0;\n\t"
            "addc.cc.u32 %3, %5, 0;\n\t"
            "madc.lo.cc.u32 %4, %9, %11, %6;\n\t"
            "madc.hi.cc.u32 %5, %9, %11, %7;\n\t"
            "madc.lo.cc.u32 %6, %10, %11, 0;\n\t"
            "madc.hi.u32 %7, %10, %11, 0;"
            : "+r"(x[0]), "+r"(x[1]), "+r"(x[2]), "+r"(x[3]), "+r"(x[4]), "+r"(x[5]), "+r"(x[6]), "+r"(x[7]),
              "+r"(y0)
            : "r"(a2), "r"(a3), "r"(b));
    } else {
        asm("add.cc.u32 %8, %8, %1;\n\t"
            "addc.cc.u32 %0, %2, 0;\n\t"
            "addc.cc.u32 %1, %3, 0;\n\t"
            "addc.cc.u32 %2, %4, 0;\n\t"
            "addc.cc.u32 %3, %5, 0;\n\t"
            "addc.cc.u32 %4, %6, 0;\n\t"
            "addc.cc.u32 %5, %7, 0;\n\t"
            "madc.lo.cc.u32 %6, %9, %10, 0;\n\t"
            "madc.hi.u32 %7, %9, %10, 0;"
            : "+r"(x[0]), "+r"(x[1]), "+r"(x[2]), "+r"(x[3]), "+r"(x[4]), "+r"(x[5]), "+r"(x[6]), "+r"(x[7]),
              "+r"(y0)
            : "r"(a3), "r"(b));
    }
#else
    const uint32_t a[4] = {a0, a1, a2, a3};
    uint64_t s = (uint64_t)y0 + x[1];
    y0 = (uint32_t)s;
    uint64_t c = s >> 32;
    for (int k = 0; k < 4; k++) {
        uint64_t add = (k < 3) ? ((((uint64_t)x[2 * k + 3]) << 32) | x[2 * k + 2]) : 0;
        unsigned __int128 w = (unsigned __int128)(k >= S ? (uint64_t)a[k] * b : 0) + add + c;
        x[2 * k] = (uint32_t)w;
        x[2 * k + 1] = (uint32_t)(w >> 32);
        c = (uint64_t)(w >> 64);
    }
#endif
}

// r = w + (z >> 32), z[0] == 0: merges the two accumulators after the last row.
PSD_DEV void merge8(uint32_t* r, const uint32_t* w, const uint32_t* z) {
#if !defined(PSD_HOST_EMU)
    asm("add.cc.u32 %0, %8, %16;\n\t"
        "addc.cc.u32 %1, %9, %17;\n\t"
        "addc.cc.u32 %2, %10, %18;\n\t"
        "addc.cc.u32 %3, %11, %19;\n\t"
        "addc.cc.u32 %4, %12, %20;\n\t"
        "addc.cc.u32 %5, %13, %21;\n\t"
        "addc.cc.u32 %6, %14, %22;\n\t"
        "addc.u32 %7, %15, 0;"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7])
        : "r"(w[0]), "r"(w[1]), "r"(w[2]), "r"(w[3]), "r"(w[4]), "r"(w[5]), "r"(w[6]), "r"(w[7]),
          "r"(z[1]), "r"(z[2]), "r"(z[3]), "r"(z[4]), "r"(z[5]), "r"(z[6]), "r"(z[7]));
#else
    uint64_t c = 0;
    for (int k = 0; k < 8; k++) {
        uint64_t s = (uint64_t)w[k] + (k < 7 ? z[k + 1] : 0) + c;
        r[k] = (uint32_t)s;
        c = s >> 32;
    }
#endif
}

// r = a + b (256-bit, no carry out expected)
PSD_DEV void add8(uint32_t* r, const uint32_t* a, const uint32_t* b) {
#if !defined(PSD_HOST_EMU)
    asm("add.cc.u32 %0, %8, %16;\n\t"
        "addc.cc.u32 %1, %9, %17;\n\t"
        "addc.cc.u32 %2, %10, %18;\n\t"
        "addc.cc.u32 %3, %11, %19;\n\t"
        "addc.cc.u32 %4, %12, %20;\n\t"
        "addc.cc.u32 %5, %13, %21;\n\t"
        "addc.cc.u32 %6, %14, %22;\n\t"
        "addc.u32 %7, %15, %23;"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7])
        : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(a[4]), "r"(a[5]), "r"(a[6]), "r"(a[7]),
          "r"(b[0]), "r"(b[1]), "r"(b[2]), "r"(b[3]), "r"(b[4]), "r"(b[5]), "r"(b[6]), "r"(b[7]));
#else
    uint64_t c = 0;
    for (int k = 0; k < 8; k++) {
        uint64_t s = (uint64_t)a[k] + b[k] + c;
        r[k] = (uint32_t)s;
        c = s >> 32;
    }
#endif
}

// r = (a >= p) ? a - p : a      (a < 2p)
template <class F>
PSD_DEV void cond_sub_p(uint32_t* r, const uint32_t* a) {
    uint32_t d[8];
    uint32_t borrow;
#if !defined(PSD_HOST_EMU)
    asm("sub.cc.u32 %0, %9, %17;\n\t"
        "subc.cc.u32 %1, %10, %18;\n\t"
        "subc.cc.u32 %2, %11, %19;\n\t"
        "subc.cc.u32 %3, %12, %20;\n\t"
        "subc.cc.u32 %4, %13, %21;\n\t"
        "subc.cc.u32 %5, %14, %22;\n\t"
        "subc.cc.u32 %6, %15, %23;\n\t"
        "subc.cc.u32 %7, %16, %24;\n\t"
        "subc.u32 %8, 0, 0;"
        : "=r"(d[0]), "=r"(d[1]), "=r"(d[2]), "=r"(d[3]), "=r"(d[4]), "=r"(d[5]), "=r"(d[6]), "=r"(d[7]),
          "=r"(borrow)
        : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(a[4]), "r"(a[5]), "r"(a[6]), "r"(a[7]),
          "n"(F::P0), "n"(F::P1), "n"(F::P2), "n"(F::P3), "n"(F::P4), "n"(F::P5), "n"(F::P6), "n"(F::P7));
#else
    const uint32_t p[8] = {F::P0, F::P1, F::P2, F::P3, F::P4, F::P5, F::P6, F::P7};
    int64_t c = 0;
    for (int k = 0; k < 8; k++) {
        int64_t s = (int64_t)a[k] - p[k] + c;
        d[k] = (uint32_t)s;
        c = s >> 32;   // 0 or -1
    }
    borrow = (uint32_t)c;
#endif
#pragma unroll
    for (int k = 0; k < 8; k++) r[k] = borrow ? a[k] : d[k];
}

// ------------------------------------------------------------------------------------------------
// The Montgomery step  acc += m * p  split over the two accumulators:
//   redc_hi: hi-aligned accumulator x (columns 1..8) += m * {p1,p3,p5,p7}, carry out added to *top
//   redc_lo: lo-aligned accumulator y (columns 0..7) += m * {p0,p2,p4,p6}, carry out added to *c8
// Generic fields use four wide multiplies each.  BLS12-381 Fr has p0 = 1 and p1 = 2^32 - 1, so
// m*p0 = m and m*p1 = (m << 32) - m: those two products become two or three adds and the step
// needs 6 wide multiplies instead of 8 (a quarter of the reduction's multiplier-pipe cycles).
// ------------------------------------------------------------------------------------------------
template <class F>
PSD_DEV void redc_hi(uint32_t* x, uint32_t m, uint32_t& top) {
#if !defined(PSD_HOST_EMU)
    if (F::kBits == 255) {
        asm("{\n\t"
            ".reg .u32 t0, t1;\n\t"
            "sub.cc.u32 t0, 0, %9;\n\t"
            "subc.u32 t1, %9, 0;\n\t"
            "add.cc.u32 %0, %0, t0;\n\t"
            "addc.cc.u32 %1, %1, t1;\n\t"
            "madc.lo.cc.u32 %2, %10, %9, %2;\n\t"
            "madc.hi.cc.u32 %3, %10, %9, %3;\n\t"
            "madc.lo.cc.u32 %4, %11, %9, %4;\n\t"
            "madc.hi.cc.u32 %5, %11, %9, %5;\n\t"
            "madc.lo.cc.u32 %6, %12, %9, %6;\n\t"
            "madc.hi.cc.u32 %7, %12, %9, %7;\n\t"
            "addc.u32 %8, %8, 0;\n\t"
            "}"
            : "+r"(x[0]), "+r"(x[1]), "+r"(x[2]), "+r"(x[3]), "+r"(x[4]), "+r"(x[5]), "+r"(x[6]), "+r"(x[7]), "+r"(top)
            : "r"(m), "r"(plimb<F, 3>()), "r"(plimb<F, 5>()), "r"(plimb<F, 7>()));
        return;
    }
#endif
    mad4(x, plimb<F, 1>(), plimb<F, 3>(), plimb<F, 5>(), plimb<F, 7>(), m, top);
}

template <class F>
PSD_DEV void redc_lo(uint32_t* y, uint32_t m, uint32_t& c8) {
#if !defined(PSD_HOST_EMU)
    if (F::kBits == 255) {
        asm("add.cc.u32 %0, %0, %9;\n\t"
            "addc.cc.u32 %1, %1, 0;\n\t"
            "madc.lo.cc.u32 %2, %10, %9, %2;\n\t"
            "madc.hi.cc.u32 %3, %10, %9, %3;\n\t"
            "madc.lo.cc.u32 %4, %11, %9, %4;\n\t"
            "madc.hi.cc.u32 %5, %11, %9, %5;\n\t"
            "madc.lo.cc.u32 %6, %12, %9, %6;\n\t"
            "madc.hi.cc.u32 %7, %12, %9, %7;\n\t"
            "addc.u32 %8, %8, 0;"
            : "+r"(y[0]), "+r"(y[1]), "+r"(y[2]), "+r"(y[3]), "+r"(y[4]), "+r"(y[5]), "+r"(y[6]), "+r"(y[7]), "+r"(c8)
            : "r"(m), "r"(plimb<F, 2>()), "r"(plimb<F, 4>()), "r"(plimb<F, 6>()));
        return;
    }
#endif
    mad4(y, plimb<F, 0>(), plimb<F, 2>(), plimb<F, 4>(), plimb<F, 6>(), m, c8);
}

// narrow window: the hi-aligned accumulator provably never carries out (see mont_mul)
template <class F>
PSD_DEV void redc_step(uint32_t* x, uint32_t* y) {
    const uint32_t m = y[0] * np0<F>();
    uint32_t unused = 0;
    if (F::kBits == 255) redc_hi<F>(x, m, unused);
    else mad4_nc(x, plimb<F, 1>(), plimb<F, 3>(), plimb<F, 5>(), plimb<F, 7>(), m);
    redc_lo<F>(y, m, x[7]);
}

// ------------------------------------------------------------------------------------------------
// Field operations on fully reduced Montgomery residues in [0,p).
// ------------------------------------------------------------------------------------------------

// r = a + b mod p   (ark-ff add_assign; used by apply_ark poseidon/mod.rs:76-80, the MDS sums 82-93
// and absorb 128,143)
template <class F>
PSD_DEV void fp_add(uint32_t* r, const uint32_t* a, const uint32_t* b) {
    uint32_t s[8];
    add8(s, a, b);          // a + b < 2p < 2^256 for both fields
    cond_sub_p<F>(r, s);
}

// One CIOS row after the first: window slides, adds a*b_i, then the Montgomery step.
// On entry x[0] == 0 (x lo-aligned, y hi-aligned); on exit y[0] == 0 with roles swapped.
template <class F>
PSD_DEV void mont_row(uint32_t* x, uint32_t* y, const uint32_t* a, uint32_t bi) {
    mad4_slide(x, y[0], a[1], a[3], a[5], a[7], bi);
    mad4(y, a[0], a[2], a[4], a[6], bi, x[7]);
    redc_step<F>(x, y);
}

// r = a * b * 2^-256 mod p, fully reduced.  Requires a < 2^255 (true for any residue of either
// field), b < 2^256.  Bound: with V the running (unshifted) accumulator value,
// V_i <= V_{i-1}/2^32 + (2^32-1)(a+p) < 2^288, so the hi-aligned accumulator (columns 1..8) never
// carries out, and the result is < a + p <= 2p: one conditional subtraction.
// 128 IMAD.WIDE.U32 + 8 IMAD = 136 integer multiplies (SURVEY.md 8(d) unit of work).
//
// REDUCE = false leaves the result in [0, a*b/2^256 + p) without the conditional subtraction (BN254
// only; see sbox_compact's LAZY mode for the bounds that make this safe).
template <class F, bool REDUCE = true>
PSD_DEV void mont_mul(uint32_t* r, const uint32_t* a, const uint32_t* b) {
    uint32_t e[8], o[8];
    mul4(e, a[0], a[2], a[4], a[6], b[0]);
    mul4(o, a[1], a[3], a[5], a[7], b[0]);
    redc_step<F>(o, e);
    mont_row<F>(e, o, a, b[1]);
    mont_row<F>(o, e, a, b[2]);
    mont_row<F>(e, o, a, b[3]);
    mont_row<F>(o, e, a, b[4]);
    mont_row<F>(e, o, a, b[5]);
    mont_row<F>(o, e, a, b[6]);
    mont_row<F>(e, o, a, b[7]);
    if (REDUCE) {
        uint32_t s[8];
        merge8(s, e, o);        // o[0] == 0 here
        cond_sub_p<F>(r, s);
    } else {
        merge8(r, e, o);
    }
}

// r = (sum_{j<N} a[j] * b[j]) * 2^-256 mod p with ONE interleaved Montgomery reduction ("lazy"
// MDS row): N*64 + 64 wide multiplies instead of N*128.  Requires every a[j] < p and
// (N+1)*p < 2^256, which keeps the running window below 2^288 (same argument as mont_mul with
// N+1 addends of size < p*2^32 per row), so the hi-aligned accumulator never carries out; the
// result is < p*(1 + N*p/2^256) < 2p: one conditional subtraction.
//   BN254 Fr: N <= 4.   BLS12-381 Fr: N = 1 only (use mont_mul).
template <class F, int N>
PSD_DEV void mont_dot(uint32_t* r, const uint32_t (*a)[8], const uint32_t (*b)[8]) {
    uint32_t e[8], o[8];
    // row 0
    mul4(e, a[0][0], a[0][2], a[0][4], a[0][6], b[0][0]);
    mul4(o, a[0][1], a[0][3], a[0][5], a[0][7], b[0][0]);
#pragma unroll
    for (int j = 1; j < N; j++) {
        mad4_nc(o, a[j][1], a[j][3], a[j][5], a[j][7], b[j][0]);
        mad4(e, a[j][0], a[j][2], a[j][4], a[j][6], b[j][0], o[7]);
    }
    {
        redc_step<F>(o, e);
    }
    // rows 1..7: x = accumulator whose column 0 is now zero, y = the other one
#pragma unroll
    for (int i = 1; i < 8; i++) {
        uint32_t* x = (i & 1) ? e : o;
        uint32_t* y = (i & 1) ? o : e;
        mad4_slide(x, y[0], a[0][1], a[0][3], a[0][5], a[0][7], b[0][i]);
        mad4(y, a[0][0], a[0][2], a[0][4], a[0][6], b[0][i], x[7]);
#pragma unroll
        for (int j = 1; j < N; j++) {
            mad4_nc(x, a[j][1], a[j][3], a[j][5], a[j][7], b[j][i]);
            mad4(y, a[j][0], a[j][2], a[j][4], a[j][6], b[j][i], x[7]);
        }
        redc_step<F>(x, y);
    }
    uint32_t s[8];
    merge8(s, e, o);        // after row 7: o[0] == 0
    cond_sub_p<F>(r, s);
}

// Largest N for which mont_dot's no-overflow condition (N+1)*p < 2^256 holds.
template <class F>
struct DotLimit {
    static constexpr int kMax = 4;                      // terms per lazily reduced group
    static constexpr bool kWide = (F::kBits != 254);    // needs the 10-limb window (mont_dot_w)
};

// One CIOS row of the squaring: only the products a_i * v_j with j >= i are added
// (v = [.., a_i at j = i, doubled limbs above]); the rest of the row is the usual Montgomery step.
template <class F, int I>
PSD_DEV void sqr_row(uint32_t* x, uint32_t* y, const uint32_t* v, uint32_t ai) {
    // odd-index entries ride the sliding (hi-aligned) chain, even-index entries the lo-aligned one
    mad4_slide_from<I / 2>(x, y[0], v[1], v[3], v[5], v[7], ai);
    mad4_from<(I + 1) / 2>(y, v[0], v[2], v[4], v[6], ai, x[7]);
    redc_step<F>(x, y);
}

// r = a^2 * 2^-256 mod p with 36 wide multiplies for the product instead of 64: each off-diagonal
// product a_i*a_j (i<j) is taken once against the doubled operand d = 2a (limb j of d carries the top
// bit of a_{j-1}; for row i that bit is masked off limb i+1, so the row adds
// a_i * (a_i*2^(32i) + 2*(a >> 32(i+1)) << 32(i+1))).  Needs 3p < 2^256 for the 9-limb window
// (row multiplicand < 2a < 2p, plus m*p), i.e. BN254 Fr; F::kSqrNineLimb says so.
// Accepts any a with 2a + p < 2^256 (BN254: a < 2.14 p), so an unreduced result of a previous
// REDUCE = false squaring can be squared again; REDUCE = false returns a value < a^2/2^256 + p.
template <class F, bool REDUCE = true>
PSD_DEV void mont_sqr(uint32_t* r, const uint32_t* a) {
    uint32_t d[8];
#if !defined(PSD_HOST_EMU)
#pragma unroll
    for (int k = 1; k < 8; k++) d[k] = __funnelshift_l(a[k - 1], a[k], 1);
#else
    for (int k = 1; k < 8; k++) d[k] = (a[k] << 1) | (a[k - 1] >> 31);
#endif
    d[0] = 0;
    uint32_t e[8], o[8];
    {   // row 0: all eight products
        const uint32_t d1 = d[1] & 0xfffffffeu;
        mul4(e, a[0], d[2], d[4], d[6], a[0]);
        mul4(o, d1, d[3], d[5], d[7], a[0]);
        redc_step<F>(o, e);
    }
    uint32_t v[8];
#define PSD_SQR_ROW(I, X, Y)                                        \
    {                                                               \
        _Pragma("unroll") for (int k = 0; k < 8; k++) v[k] = d[k];  \
        v[I] = a[I];                                                \
        if (I < 7) v[I + 1] = d[I + 1] & 0xfffffffeu;               \
        sqr_row<F, I>(X, Y, v, a[I]);                               \
    }
    PSD_SQR_ROW(1, e, o)
    PSD_SQR_ROW(2, o, e)
    PSD_SQR_ROW(3, e, o)
    PSD_SQR_ROW(4, o, e)
    PSD_SQR_ROW(5, e, o)
    PSD_SQR_ROW(6, o, e)
    PSD_SQR_ROW(7, e, o)
#undef PSD_SQR_ROW
    if (REDUCE) {
        uint32_t s[8];
        merge8(s, e, o);
        cond_sub_p<F>(r, s);
    } else {
        merge8(r, e, o);
    }
}

// ------------------------------------------------------------------------------------------------
// Unreduced products (variant 7: the reduction is a product by constants and goes to the tensor cores).
// Same two sliding accumulators as mont_mul / mont_sqr without the Montgomery steps: after row i the lowest
// column of the lo-aligned accumulator is limb i of the product.  Both accumulators hold parts of an exact
// partial sum below 2^(256 + 32 (i + 1)), so no chain carries out.
// ------------------------------------------------------------------------------------------------

// out[0..15] = a * b   (64 wide multiplies).  (The fixed-accumulator form of sqr_wide below needs as many carry adds here
// as this sliding window, 23 against 22, and measured 0.3 % slower on C2.)
PSD_DEV void mul_wide(uint32_t* out, const uint32_t* a, const uint32_t* b) {
    uint32_t e[8], o[8];
    mul4(e, a[0], a[2], a[4], a[6], b[0]);
    mul4(o, a[1], a[3], a[5], a[7], b[0]);
    out[0] = e[0];
#define PSD_WIDE_ROW(I, X, Y)                                  \
    mad4_slide(X, Y[0], a[1], a[3], a[5], a[7], b[I]);         \
    mad4(Y, a[0], a[2], a[4], a[6], b[I], X[7]);               \
    out[I] = Y[0];
    PSD_WIDE_ROW(1, e, o)
    PSD_WIDE_ROW(2, o, e)
    PSD_WIDE_ROW(3, e, o)
    PSD_WIDE_ROW(4, o, e)
    PSD_WIDE_ROW(5, e, o)
    PSD_WIDE_ROW(6, o, e)
    PSD_WIDE_ROW(7, e, o)
#undef PSD_WIDE_ROW
    merge8(out + 8, e, o);   // o[0] (= out[7]) is skipped by merge8
}

// ---- building blocks of the unreduced squaring: one carry chain over consecutive 64-bit pairs of an accumulator -------
// N pairs that already hold a value are accumulated into (p += x_k * b), then the chain ends in one of three ways:
//   _tail:  the carry out goes to the next limb, which nothing has written yet (t = carry);
//   _half:  one more product lands on a pair whose low limb holds a value and whose high limb is new
//           (product + low limb + carry < 2^64: no carry out);
//   _fresh: one more product lands on a pair nothing has written yet.
PSD_DEV void sq_pair_c(uint32_t& lo, uint32_t& hi, uint32_t x, uint32_t b, uint32_t& carry) {   // host model of one pair
    const unsigned __int128 w = (unsigned __int128)((uint64_t)x * b) + (((uint64_t)hi << 32) | lo) + carry;
    lo = (uint32_t)w;
    hi = (uint32_t)(w >> 32);
    carry = (uint32_t)(w >> 64);
}
PSD_DEV void sq_acc3_fresh(uint32_t& p0, uint32_t& p1, uint32_t& p2, uint32_t& p3, uint32_t& p4, uint32_t& p5, uint32_t& f0,
                           uint32_t& f1, uint32_t x0, uint32_t x1, uint32_t x2, uint32_t x3, uint32_t b) {
#if !defined(PSD_HOST_EMU)
    asm("mad.lo.cc.u32 %0, %8, %12, %0;\n\t"
        "madc.hi.cc.u32 %1, %8, %12, %1;\n\t"
        "madc.lo.cc.u32 %2, %9, %12, %2;\n\t"
        "madc.hi.cc.u32 %3, %9, %12, %3;\n\t"
        "madc.lo.cc.u32 %4, %10, %12, %4;\n\t"
        "madc.hi.cc.u32 %5, %10, %12, %5;\n\t"
        "madc.lo.cc.u32 %6, %11, %12, 0;\n\t"
        "madc.hi.u32 %7, %11, %12, 0;"
        : "+r"(p0), "+r"(p1), "+r"(p2), "+r"(p3), "+r"(p4), "+r"(p5), "=&r"(f0), "=&r"(f1)
        : "r"(x0), "r"(x1), "r"(x2), "r"(x3), "r"(b));
#else
    uint32_t c = 0;
    f0 = f1 = 0;
    sq_pair_c(p0, p1, x0, b, c);
    sq_pair_c(p2, p3, x1, b, c);
    sq_pair_c(p4, p5, x2, b, c);
    sq_pair_c(f0, f1, x3, b, c);
#endif
}
PSD_DEV void sq_acc3_tail(uint32_t& p0, uint32_t& p1, uint32_t& p2, uint32_t& p3, uint32_t& p4, uint32_t& p5, uint32_t& t,
                          uint32_t x0, uint32_t x1, uint32_t x2, uint32_t b) {
#if !defined(PSD_HOST_EMU)
    asm("mad.lo.cc.u32 %0, %7, %10, %0;\n\t"
        "madc.hi.cc.u32 %1, %7, %10, %1;\n\t"
        "madc.lo.cc.u32 %2, %8, %10, %2;\n\t"
        "madc.hi.cc.u32 %3, %8, %10, %3;\n\t"
        "madc.lo.cc.u32 %4, %9, %10, %4;\n\t"
        "madc.hi.cc.u32 %5, %9, %10, %5;\n\t"
        "addc.u32 %6, 0, 0;"
        : "+r"(p0), "+r"(p1), "+r"(p2), "+r"(p3), "+r"(p4), "+r"(p5), "=&r"(t)
        : "r"(x0), "r"(x1), "r"(x2), "r"(b));
#else
    uint32_t c = 0;
    sq_pair_c(p0, p1, x0, b, c);
    sq_pair_c(p2, p3, x1, b, c);
    sq_pair_c(p4, p5, x2, b, c);
    t = c;
#endif
}
PSD_DEV void sq_acc2_tail(uint32_t& p0, uint32_t& p1, uint32_t& p2, uint32_t& p3, uint32_t& t, uint32_t x0, uint32_t x1,
                          uint32_t b) {
#if !defined(PSD_HOST_EMU)
    asm("mad.lo.cc.u32 %0, %5, %7, %0;\n\t"
        "madc.hi.cc.u32 %1, %5, %7, %1;\n\t"
        "madc.lo.cc.u32 %2, %6, %7, %2;\n\t"
        "madc.hi.cc.u32 %3, %6, %7, %3;\n\t"
        "addc.u32 %4, 0, 0;"
        : "+r"(p0), "+r"(p1), "+r"(p2), "+r"(p3), "=&r"(t)
        : "r"(x0), "r"(x1), "r"(b));
#else
    uint32_t c = 0;
    sq_pair_c(p0, p1, x0, b, c);
    sq_pair_c(p2, p3, x1, b, c);
    t = c;
#endif
}
PSD_DEV void sq_acc1_tail(uint32_t& p0, uint32_t& p1, uint32_t& t, uint32_t x0, uint32_t b) {
#if !defined(PSD_HOST_EMU)
    asm("mad.lo.cc.u32 %0, %3, %4, %0;\n\t"
        "madc.hi.cc.u32 %1, %3, %4, %1;\n\t"
        "addc.u32 %2, 0, 0;"
        : "+r"(p0), "+r"(p1), "=&r"(t)
        : "r"(x0), "r"(b));
#else
    uint32_t c = 0;
    sq_pair_c(p0, p1, x0, b, c);
    t = c;
#endif
}
PSD_DEV void sq_acc2_half(uint32_t& p0, uint32_t& p1, uint32_t& p2, uint32_t& p3, uint32_t& hl, uint32_t& hh, uint32_t x0,
                          uint32_t x1, uint32_t x2, uint32_t b) {
#if !defined(PSD_HOST_EMU)
    asm("mad.lo.cc.u32 %0, %6, %9, %0;\n\t"
        "madc.hi.cc.u32 %1, %6, %9, %1;\n\t"
        "madc.lo.cc.u32 %2, %7, %9, %2;\n\t"
        "madc.hi.cc.u32 %3, %7, %9, %3;\n\t"
        "madc.lo.cc.u32 %4, %8, %9, %4;\n\t"
        "madc.hi.u32 %5, %8, %9, 0;"
        : "+r"(p0), "+r"(p1), "+r"(p2), "+r"(p3), "+r"(hl), "=&r"(hh)
        : "r"(x0), "r"(x1), "r"(x2), "r"(b));
#else
    uint32_t c = 0;
    hh = 0;
    sq_pair_c(p0, p1, x0, b, c);
    sq_pair_c(p2, p3, x1, b, c);
    sq_pair_c(hl, hh, x2, b, c);
#endif
}
PSD_DEV void sq_acc1_half(uint32_t& p0, uint32_t& p1, uint32_t& hl, uint32_t& hh, uint32_t x0, uint32_t x1, uint32_t b) {
#if !defined(PSD_HOST_EMU)
    asm("mad.lo.cc.u32 %0, %4, %6, %0;\n\t"
        "madc.hi.cc.u32 %1, %4, %6, %1;\n\t"
        "madc.lo.cc.u32 %2, %5, %6, %2;\n\t"
        "madc.hi.u32 %3, %5, %6, 0;"
        : "+r"(p0), "+r"(p1), "+r"(hl), "=&r"(hh)
        : "r"(x0), "r"(x1), "r"(b));
#else
    uint32_t c = 0;
    hh = 0;
    sq_pair_c(p0, p1, x0, b, c);
    sq_pair_c(hl, hh, x1, b, c);
#endif
}
PSD_DEV void sq_acc0_half(uint32_t& hl, uint32_t& hh, uint32_t x0, uint32_t b) {
#if !defined(PSD_HOST_EMU)
    asm("mad.lo.cc.u32 %0, %2, %3, %0;\n\t"
        "madc.hi.u32 %1, %2, %3, 0;"
        : "+r"(hl), "=&r"(hh)
        : "r"(x0), "r"(b));
#else
    uint32_t c = 0;
    hh = 0;
    sq_pair_c(hl, hh, x0, b, c);
#endif
}

// r[1..15] += s[1..14] (the odd accumulator onto the even one), in place
PSD_DEV void merge16(uint32_t* E, const uint32_t* O) {
#if !defined(PSD_HOST_EMU)
    asm("add.cc.u32 %0, %0, %15;\n\t"
        "addc.cc.u32 %1, %1, %16;\n\t"
        "addc.cc.u32 %2, %2, %17;\n\t"
        "addc.cc.u32 %3, %3, %18;\n\t"
        "addc.cc.u32 %4, %4, %19;\n\t"
        "addc.cc.u32 %5, %5, %20;\n\t"
        "addc.cc.u32 %6, %6, %21;\n\t"
        "addc.cc.u32 %7, %7, %22;\n\t"
        "addc.cc.u32 %8, %8, %23;\n\t"
        "addc.cc.u32 %9, %9, %24;\n\t"
        "addc.cc.u32 %10, %10, %25;\n\t"
        "addc.cc.u32 %11, %11, %26;\n\t"
        "addc.cc.u32 %12, %12, %27;\n\t"
        "addc.cc.u32 %13, %13, %28;\n\t"
        "addc.u32 %14, %14, 0;"
        : "+r"(E[1]), "+r"(E[2]), "+r"(E[3]), "+r"(E[4]), "+r"(E[5]), "+r"(E[6]), "+r"(E[7]), "+r"(E[8]), "+r"(E[9]),
          "+r"(E[10]), "+r"(E[11]), "+r"(E[12]), "+r"(E[13]), "+r"(E[14]), "+r"(E[15])
        : "r"(O[1]), "r"(O[2]), "r"(O[3]), "r"(O[4]), "r"(O[5]), "r"(O[6]), "r"(O[7]), "r"(O[8]), "r"(O[9]), "r"(O[10]),
          "r"(O[11]), "r"(O[12]), "r"(O[13]), "r"(O[14]));
#else
    uint64_t c = 0;
    for (int k = 1; k < 16; k++) {
        const uint64_t t = (uint64_t)E[k] + (k < 15 ? O[k] : 0) + c;
        E[k] = (uint32_t)t;
        c = t >> 32;
    }
#endif
}

// the same through the sliding two-accumulator window of mont_sqr (46 carry adds; kept for comparison, -DPSD_SQR_SLIDE)
PSD_DEV void sqr_wide_slide(uint32_t* out, const uint32_t* a) {
    uint32_t d[8];
#if !defined(PSD_HOST_EMU)
#pragma unroll
    for (int k = 1; k < 8; k++) d[k] = __funnelshift_l(a[k - 1], a[k], 1);
#else
    for (int k = 1; k < 8; k++) d[k] = (a[k] << 1) | (a[k - 1] >> 31);
#endif
    d[0] = 0;
    uint32_t e[8], o[8];
    {
        const uint32_t d1 = d[1] & 0xfffffffeu;
        mul4(e, a[0], d[2], d[4], d[6], a[0]);
        mul4(o, d1, d[3], d[5], d[7], a[0]);
        out[0] = e[0];
    }
    uint32_t v[8];
#define PSD_SQRW_ROW(I, X, Y)                                                    \
    {                                                                            \
        _Pragma("unroll") for (int k = 0; k < 8; k++) v[k] = d[k];               \
        v[I] = a[I];                                                             \
        if (I < 7) v[I + 1] = d[I + 1] & 0xfffffffeu;                            \
        mad4_slide_from<I / 2>(X, Y[0], v[1], v[3], v[5], v[7], a[I]);           \
        mad4_from<(I + 1) / 2>(Y, v[0], v[2], v[4], v[6], a[I], X[7]);           \
        out[I] = Y[0];                                                           \
    }
    PSD_SQRW_ROW(1, e, o)
    PSD_SQRW_ROW(2, o, e)
    PSD_SQRW_ROW(3, e, o)
    PSD_SQRW_ROW(4, o, e)
    PSD_SQRW_ROW(5, e, o)
    PSD_SQRW_ROW(6, o, e)
    PSD_SQRW_ROW(7, e, o)
#undef PSD_SQRW_ROW
    merge8(out + 8, e, o);
}

// out[0..15] = a^2, a < 2^255   (36 wide multiplies, 21 carry adds).
// Each off-diagonal limb product is taken once against the doubled operand d = 2a, as in mont_sqr: row i adds
// a_i * (a_i 2^(32 i) + (limbs i+1.. of d, the bit that a_i contributed masked off)) 2^(32 i).  Two FIXED 16-limb
// accumulators instead of mont_sqr's sliding window: E takes the products at even columns, O those at odd columns, so
// every chain runs over consecutive 64-bit register pairs and nothing is shifted (the sliding window spends 24 adds per
// squaring on carrying finished low columns along).  Which limbs a chain meets for the first time is known statically
// per row (the _fresh/_half/_tail endings above); the result is E + O in one 15-add chain.
PSD_DEV void sqr_wide(uint32_t* out, const uint32_t* a) {
#if defined(PSD_SQR_SLIDE)
    sqr_wide_slide(out, a);
#else
    uint32_t d[8], m[8];   // d = 2a limbs 1..7; m[k] = d[k] with the bit of a[k-1] masked off
#if !defined(PSD_HOST_EMU)
#pragma unroll
    for (int k = 1; k < 8; k++) d[k] = __funnelshift_l(a[k - 1], a[k], 1);
#else
    for (int k = 1; k < 8; k++) d[k] = (a[k] << 1) | (a[k - 1] >> 31);
#endif
#pragma unroll
    for (int k = 1; k < 8; k++) m[k] = d[k] & 0xfffffffeu;
    uint32_t E[16], O[16];   // O[c] = column c of the odd accumulator (columns 1..14)
    // row 0: eight products, every pair new
    mul4(E, a[0], d[2], d[4], d[6], a[0]);
    mul4(O + 1, m[1], d[3], d[5], d[7], a[0]);
    // row 1: E columns 2,4,6,8 (a1 a1, d3, d5, d7); O columns 3,5,7 (m2, d4, d6)
    sq_acc3_fresh(E[2], E[3], E[4], E[5], E[6], E[7], E[8], E[9], a[1], d[3], d[5], d[7], a[1]);
    sq_acc3_tail(O[3], O[4], O[5], O[6], O[7], O[8], O[9], m[2], d[4], d[6], a[1]);
    // row 2: E columns 4,6,8 (a2, d4, d6); O columns 5,7,9 (m3, d5, d7)
    sq_acc3_tail(E[4], E[5], E[6], E[7], E[8], E[9], E[10], a[2], d[4], d[6], a[2]);
    sq_acc2_half(O[5], O[6], O[7], O[8], O[9], O[10], m[3], d[5], d[7], a[2]);
    // row 3: E columns 6,8,10 (a3, d5, d7); O columns 7,9 (m4, d6)
    sq_acc2_half(E[6], E[7], E[8], E[9], E[10], E[11], a[3], d[5], d[7], a[3]);
    sq_acc2_tail(O[7], O[8], O[9], O[10], O[11], m[4], d[6], a[3]);
    // row 4: E columns 8,10 (a4, d6); O columns 9,11 (m5, d7)
    sq_acc2_tail(E[8], E[9], E[10], E[11], E[12], a[4], d[6], a[4]);
    sq_acc1_half(O[9], O[10], O[11], O[12], m[5], d[7], a[4]);
    // row 5: E columns 10,12 (a5, d7); O column 11 (m6)
    sq_acc1_half(E[10], E[11], E[12], E[13], a[5], d[7], a[5]);
    sq_acc1_tail(O[11], O[12], O[13], m[6], a[5]);
    // row 6: E column 12 (a6); O column 13 (m7)
    sq_acc1_tail(E[12], E[13], E[14], a[6], a[6]);
    sq_acc0_half(O[13], O[14], m[7], a[6]);
    // row 7: E column 14 (a7)
    sq_acc0_half(E[14], E[15], a[7], a[7]);
    merge16(E, O);
#pragma unroll
    for (int k = 0; k < 16; k++) out[k] = E[k];
#endif
}

// ------------------------------------------------------------------------------------------------
// Wide-window variants.  For BLS12-381 Fr (p ~ 0.45 * 2^256) the running window of a lazily reduced
// dot product or of the squaring can exceed 9 limbs, so a tenth limb `top` (column 9) collects the
// carries out of column 8 and re-enters at the top pair when the window slides.  The result can be
// up to 3p (a 257-bit value); reduce_3p brings it back to [0,p).
// ------------------------------------------------------------------------------------------------
// x7 += c; top += carry
PSD_DEV void add_carry2(uint32_t& x7, uint32_t& top, uint32_t c) {
#if !defined(PSD_HOST_EMU)
    asm("add.cc.u32 %0, %0, %2;\n\taddc.u32 %1, %1, 0;" : "+r"(x7), "+r"(top) : "r"(c));
#else
    uint64_t s = (uint64_t)x7 + c;
    x7 = (uint32_t)s;
    top += (uint32_t)(s >> 32);
#endif
}
// lo-aligned chain (columns 0..7): carry -> column 8 (x7) -> column 9 (top)
template <int S>
PSD_DEV void mad4w_lo(uint32_t* y, uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b, uint32_t& x7,
                      uint32_t& top) {
    if (S >= 4) return;
    uint32_t c = 0;
    mad4_from<S>(y, a0, a1, a2, a3, b, c);
    add_carry2(x7, top, c);
}
// window slide: as mad4_slide_from, then the old column 9 re-enters at column 8 and the new
// column 9 is the carry out of it
template <int S>
PSD_DEV void mad4w_slide(uint32_t* x, uint32_t& y0, uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b,
                         uint32_t& top) {
    mad4_slide_from<S>(x, y0, a0, a1, a2, a3, b);
    const uint32_t old = top;
    top = 0;
    add_carry2(x[7], top, old);
}
// Montgomery step on a wide window
template <class F>
PSD_DEV void redc_step_w(uint32_t* x, uint32_t* y, uint32_t& top) {
    const uint32_t m = y[0] * np0<F>();
    redc_hi<F>(x, m, top);
    uint32_t c = 0;
    redc_lo<F>(y, m, c);
    add_carry2(x[7], top, c);
}
// r = (w + (z >> 32) + top * 2^256) mod p for a value below 3p
template <class F>
PSD_DEV void reduce_3p(uint32_t* r, const uint32_t* w, const uint32_t* z, uint32_t top) {
    uint32_t s[8], c8;
#if !defined(PSD_HOST_EMU)
    asm("add.cc.u32 %0, %9, %17;\n\t"
        "addc.cc.u32 %1, %10, %18;\n\t"
        "addc.cc.u32 %2, %11, %19;\n\t"
        "addc.cc.u32 %3, %12, %20;\n\t"
        "addc.cc.u32 %4, %13, %21;\n\t"
        "addc.cc.u32 %5, %14, %22;\n\t"
        "addc.cc.u32 %6, %15, %23;\n\t"
        "addc.cc.u32 %7, %16, 0;\n\t"
        "addc.u32 %8, %24, 0;"
        : "=r"(s[0]), "=r"(s[1]), "=r"(s[2]), "=r"(s[3]), "=r"(s[4]), "=r"(s[5]), "=r"(s[6]), "=r"(s[7]), "=r"(c8)
        : "r"(w[0]), "r"(w[1]), "r"(w[2]), "r"(w[3]), "r"(w[4]), "r"(w[5]), "r"(w[6]), "r"(w[7]),
          "r"(z[1]), "r"(z[2]), "r"(z[3]), "r"(z[4]), "r"(z[5]), "r"(z[6]), "r"(z[7]), "r"(top));
#else
    uint64_t c = 0;
    for (int k = 0; k < 8; k++) {
        uint64_t t = (uint64_t)w[k] + (k < 7 ? z[k + 1] : 0) + c;
        s[k] = (uint32_t)t;
        c = t >> 32;
    }
    c8 = top + (uint32_t)c;
#endif
    // step 1: v = c8*2^256 + s < 3p.  d = s - p mod 2^256; take d if c8 is set (then v - p < 2^256) or s >= p
    uint32_t d[8], borrow;
#if !defined(PSD_HOST_EMU)
    asm("sub.cc.u32 %0, %9, %17;\n\t"
        "subc.cc.u32 %1, %10, %18;\n\t"
        "subc.cc.u32 %2, %11, %19;\n\t"
        "subc.cc.u32 %3, %12, %20;\n\t"
        "subc.cc.u32 %4, %13, %21;\n\t"
        "subc.cc.u32 %5, %14, %22;\n\t"
        "subc.cc.u32 %6, %15, %23;\n\t"
        "subc.cc.u32 %7, %16, %24;\n\t"
        "subc.u32 %8, 0, 0;"
        : "=r"(d[0]), "=r"(d[1]), "=r"(d[2]), "=r"(d[3]), "=r"(d[4]), "=r"(d[5]), "=r"(d[6]), "=r"(d[7]), "=r"(borrow)
        : "r"(s[0]), "r"(s[1]), "r"(s[2]), "r"(s[3]), "r"(s[4]), "r"(s[5]), "r"(s[6]), "r"(s[7]),
          "r"(plimb<F, 0>()), "r"(plimb<F, 1>()), "r"(plimb<F, 2>()), "r"(plimb<F, 3>()), "r"(plimb<F, 4>()),
          "r"(plimb<F, 5>()), "r"(plimb<F, 6>()), "r"(plimb<F, 7>()));
#else
    const uint32_t p[8] = {F::P0, F::P1, F::P2, F::P3, F::P4, F::P5, F::P6, F::P7};
    int64_t bc = 0;
    for (int k = 0; k < 8; k++) {
        int64_t t = (int64_t)s[k] - p[k] + bc;
        d[k] = (uint32_t)t;
        bc = t >> 32;
    }
    borrow = (uint32_t)bc;
#endif
    const bool take = (c8 != 0) || (borrow == 0);
    uint32_t v[8];
#pragma unroll
    for (int k = 0; k < 8; k++) v[k] = take ? d[k] : s[k];
    // step 2: v < 2p
    cond_sub_p<F>(r, v);
}

// Lazily reduced dot product with the wide window: any field here, N <= 4 (result < 3p).
template <class F, int N>
PSD_DEV void mont_dot_w(uint32_t* r, const uint32_t (*a)[8], const uint32_t (*b)[8]) {
    uint32_t e[8], o[8], top = 0;
    mul4(e, a[0][0], a[0][2], a[0][4], a[0][6], b[0][0]);
    mul4(o, a[0][1], a[0][3], a[0][5], a[0][7], b[0][0]);
#pragma unroll
    for (int j = 1; j < N; j++) {
        mad4(o, a[j][1], a[j][3], a[j][5], a[j][7], b[j][0], top);
        mad4w_lo<0>(e, a[j][0], a[j][2], a[j][4], a[j][6], b[j][0], o[7], top);
    }
    redc_step_w<F>(o, e, top);
#pragma unroll
    for (int i = 1; i < 8; i++) {
        uint32_t* x = (i & 1) ? e : o;
        uint32_t* y = (i & 1) ? o : e;
        mad4w_slide<0>(x, y[0], a[0][1], a[0][3], a[0][5], a[0][7], b[0][i], top);
        mad4w_lo<0>(y, a[0][0], a[0][2], a[0][4], a[0][6], b[0][i], x[7], top);
#pragma unroll
        for (int j = 1; j < N; j++) {
            mad4(x, a[j][1], a[j][3], a[j][5], a[j][7], b[j][i], top);
            mad4w_lo<0>(y, a[j][0], a[j][2], a[j][4], a[j][6], b[j][i], x[7], top);
        }
        redc_step_w<F>(x, y, top);
    }
    reduce_3p<F>(r, e, o, top);
}

template <class F, int I>
PSD_DEV void sqr_row_w(uint32_t* x, uint32_t* y, const uint32_t* v, uint32_t ai, uint32_t& top) {
    mad4w_slide<I / 2>(x, y[0], v[1], v[3], v[5], v[7], ai, top);
    mad4w_lo<(I + 1) / 2>(y, v[0], v[2], v[4], v[6], ai, x[7], top);
    redc_step_w<F>(x, y, top);
}

// mont_sqr with the wide window (same product schedule): any field with p < 2^255.
template <class F>
PSD_DEV void mont_sqr_w(uint32_t* r, const uint32_t* a) {
    uint32_t d[8];
#if !defined(PSD_HOST_EMU)
#pragma unroll
    for (int k = 1; k < 8; k++) d[k] = __funnelshift_l(a[k - 1], a[k], 1);
#else
    for (int k = 1; k < 8; k++) d[k] = (a[k] << 1) | (a[k - 1] >> 31);
#endif
    d[0] = 0;
    uint32_t e[8], o[8], top = 0;
    {
        const uint32_t d1 = d[1] & 0xfffffffeu;
        mul4(e, a[0], d[2], d[4], d[6], a[0]);
        mul4(o, d1, d[3], d[5], d[7], a[0]);
        redc_step_w<F>(o, e, top);
    }
    uint32_t v[8];
#define PSD_SQR_ROW_W(I, X, Y)                                      \
    {                                                               \
        _Pragma("unroll") for (int k = 0; k < 8; k++) v[k] = d[k];  \
        v[I] = a[I];                                                \
        if (I < 7) v[I + 1] = d[I + 1] & 0xfffffffeu;               \
        sqr_row_w<F, I>(X, Y, v, a[I], top);                        \
    }
    PSD_SQR_ROW_W(1, e, o)
    PSD_SQR_ROW_W(2, o, e)
    PSD_SQR_ROW_W(3, e, o)
    PSD_SQR_ROW_W(4, o, e)
    PSD_SQR_ROW_W(5, e, o)
    PSD_SQR_ROW_W(6, o, e)
    PSD_SQR_ROW_W(7, e, o)
#undef PSD_SQR_ROW_W
    reduce_3p<F>(r, e, o, top);
}

// ------------------------------------------------------------------------------------------------
// Multiplication by CONSTANTS without a full Montgomery reduction ("digit tables").
//
// For a constant c the host precomputes  tab[d] = c * 2^(64 d) * 2^96 mod p  (d = 0..3, canonical).
// For any a = sum_d D_d 2^(64 d) (64-bit digits D_d = (lo, hi) limbs):
//     a * c * 2^96  ==  sum_d D_d * tab[d]   (mod p)
// The right-hand side is 4 products of a 64-bit digit with a 256-bit table entry -- 64 wide
// multiplies, all accumulated at FIXED alignment into an 11-limb accumulator (no per-row reduction, no
// sliding) -- and the sum S < N * 2^66 * p of N such products needs only THREE Montgomery rows to
// divide the 2^96 back out: S * 2^-96 + (<p*(1+2^-31)) < 2p.  So a constant product costs
// 64 + 24 wide multiplies instead of 128, a sum of N constant products 64 N + 24 instead of 64 N + 64,
// for ANY N and either field.  If `a` is a Montgomery residue, so is the result.
//
// Accumulator: e[10] = columns 0..9 (pairs at even columns), o[8] = columns 1..8 (pairs at odd
// columns), top = column 10.
// ------------------------------------------------------------------------------------------------
// d[0..7] (four pairs) += {a0..a3} * b; the carry out ripples into r0, r1, top
PSD_DEV void mad4_rip3(uint32_t* d, uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b, uint32_t& r0,
                       uint32_t& r1, uint32_t& top) {
#if !defined(PSD_HOST_EMU)
    asm("mad.lo.cc.u32 %0, %11, %15, %0;\n\t"
        "madc.hi.cc.u32 %1, %11, %15, %1;\n\t"
        "madc.lo.cc.u32 %2, %12, %15, %2;\n\t"
        "madc.hi.cc.u32 %3, %12, %15, %3;\n\t"
        "madc.lo.cc.u32 %4, %13, %15, %4;\n\t"
        "madc.hi.cc.u32 %5, %13, %15, %5;\n\t"
        "madc.lo.cc.u32 %6, %14, %15, %6;\n\t"
        "madc.hi.cc.u32 %7, %14, %15, %7;\n\t"
        "addc.cc.u32 %8, %8, 0;\n\t"
        "addc.cc.u32 %9, %9, 0;\n\t"
        "addc.u32 %10, %10, 0;"
        : "+r"(d[0]), "+r"(d[1]), "+r"(d[2]), "+r"(d[3]), "+r"(d[4]), "+r"(d[5]), "+r"(d[6]), "+r"(d[7]), "+r"(r0),
          "+r"(r1), "+r"(top)
        : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b));
#else
    uint32_t c = 0;
    mad4(d, a0, a1, a2, a3, b, c);
    uint64_t s = (uint64_t)r0 + c;
    r0 = (uint32_t)s;
    s = (uint64_t)r1 + (s >> 32);
    r1 = (uint32_t)s;
    top += (uint32_t)(s >> 32);
#endif
}
// same, carry ripples into r1, top
PSD_DEV void mad4_rip2(uint32_t* d, uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b, uint32_t& r1,
                       uint32_t& top) {
#if !defined(PSD_HOST_EMU)
    asm("mad.lo.cc.u32 %0, %10, %14, %0;\n\t"
        "madc.hi.cc.u32 %1, %10, %14, %1;\n\t"
        "madc.lo.cc.u32 %2, %11, %14, %2;\n\t"
        "madc.hi.cc.u32 %3, %11, %14, %3;\n\t"
        "madc.lo.cc.u32 %4, %12, %14, %4;\n\t"
        "madc.hi.cc.u32 %5, %12, %14, %5;\n\t"
        "madc.lo.cc.u32 %6, %13, %14, %6;\n\t"
        "madc.hi.cc.u32 %7, %13, %14, %7;\n\t"
        "addc.cc.u32 %8, %8, 0;\n\t"
        "addc.u32 %9, %9, 0;"
        : "+r"(d[0]), "+r"(d[1]), "+r"(d[2]), "+r"(d[3]), "+r"(d[4]), "+r"(d[5]), "+r"(d[6]), "+r"(d[7]), "+r"(r1),
          "+r"(top)
        : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b));
#else
    uint32_t c = 0;
    mad4(d, a0, a1, a2, a3, b, c);
    uint64_t s = (uint64_t)r1 + c;
    r1 = (uint32_t)s;
    top += (uint32_t)(s >> 32);
#endif
}

// acc += (lo + hi*2^32) * t[0..7].  Carries out of the four chains are only COUNTED (c8, c9, c10 =
// carries into columns 8, 9, 10); cacc_finish folds the counts in once.
PSD_DEV void cacc_digit(uint32_t* e, uint32_t* o, uint32_t& c8, uint32_t& c9, uint32_t& c10, uint32_t lo, uint32_t hi,
                        const uint32_t* t) {
    mad4(e, t[0], t[2], t[4], t[6], lo, c8);        // columns (0,1)..(6,7)
    mad4(o, t[1], t[3], t[5], t[7], lo, c9);        // columns (1,2)..(7,8)
    mad4(o, t[0], t[2], t[4], t[6], hi, c9);        // columns (1,2)..(7,8)
    mad4(e + 2, t[1], t[3], t[5], t[7], hi, c10);   // columns (2,3)..(8,9)
}

// x7 += c; t9 += carry; t10 += carry
PSD_DEV void add_carry3(uint32_t& x7, uint32_t& t9, uint32_t& t10, uint32_t c) {
#if !defined(PSD_HOST_EMU)
    asm("add.cc.u32 %0, %0, %3;\n\taddc.cc.u32 %1, %1, 0;\n\taddc.u32 %2, %2, 0;" : "+r"(x7), "+r"(t9), "+r"(t10) : "r"(c));
#else
    uint64_t s = (uint64_t)x7 + c;
    x7 = (uint32_t)s;
    s = (uint64_t)t9 + (s >> 32);
    t9 = (uint32_t)s;
    t10 += (uint32_t)(s >> 32);
#endif
}

// One Montgomery row on an 11-column value: hi = hi-aligned (columns 1..8), lo = lo-aligned (0..7),
// t9/t10 = columns 9/10.  Afterwards lo[0] == 0.
template <class F>
PSD_DEV void redc11(uint32_t* hi, uint32_t* lo, uint32_t& t9, uint32_t& t10) {
    const uint32_t m = lo[0] * np0<F>();
    uint32_t c = 0;
    redc_hi<F>(hi, m, c);
    add_carry2(t9, t10, c);
    c = 0;
    redc_lo<F>(lo, m, c);
    add_carry3(hi[7], t9, t10, c);
}

// The window moves down one column: x (lo-aligned, x[0] == 0) becomes the hi-aligned accumulator,
// y0 is column 0 of the other one; columns 9/10 shift down with it.
PSD_DEV void slide11(uint32_t* x, uint32_t& y0, uint32_t& t9, uint32_t& t10) {
#if !defined(PSD_HOST_EMU)
    asm("add.cc.u32 %8, %8, %1;\n\t"
        "addc.cc.u32 %0, %2, 0;\n\t"
        "addc.cc.u32 %1, %3, 0;\n\t"
        "addc.cc.u32 %2, %4, 0;\n\t"
        "addc.cc.u32 %3, %5, 0;\n\t"
        "addc.cc.u32 %4, %6, 0;\n\t"
        "addc.cc.u32 %5, %7, 0;\n\t"
        "addc.cc.u32 %6, 0, 0;\n\t"
        "addc.cc.u32 %7, %9, 0;\n\t"
        "addc.u32 %9, %10, 0;"
        : "+r"(x[0]), "+r"(x[1]), "+r"(x[2]), "+r"(x[3]), "+r"(x[4]), "+r"(x[5]), "+r"(x[6]), "+r"(x[7]), "+r"(y0),
          "+r"(t9)
        : "r"(t10));
#else
    uint64_t s = (uint64_t)y0 + x[1];
    y0 = (uint32_t)s;
    uint64_t c = s >> 32;
    for (int k = 0; k < 6; k++) {
        s = (uint64_t)x[k + 2] + c;
        x[k] = (uint32_t)s;
        c = s >> 32;
    }
    x[6] = (uint32_t)c;
    x[7] = t9;
    t9 = t10;
#endif
    t10 = 0;
}

// r = S * 2^-96 mod p for the accumulated sum S (< 2^70 * p), fully reduced.
template <class F>
PSD_DEV void cacc_finish(uint32_t* r, const uint32_t* e, const uint32_t* o, uint32_t c8, uint32_t c9, uint32_t c10) {
    uint32_t y[8], x[8], t9, t10;
#pragma unroll
    for (int k = 0; k < 8; k++) y[k] = e[k];
#pragma unroll
    for (int k = 0; k < 7; k++) x[k] = o[k];
    // column 8 = o[7] + e[8] + c8, column 9 = e[9] + c9 + carry, column 10 = c10 + carry
#if !defined(PSD_HOST_EMU)
    asm("add.cc.u32 %0, %3, %4;\n\t"
        "addc.cc.u32 %1, %5, %7;\n\t"
        "addc.u32 %2, %8, 0;\n\t"
        "add.cc.u32 %0, %0, %6;\n\t"
        "addc.cc.u32 %1, %1, 0;\n\t"
        "addc.u32 %2, %2, 0;"
        : "=&r"(x[7]), "=&r"(t9), "=&r"(t10)
        : "r"(o[7]), "r"(e[8]), "r"(e[9]), "r"(c8), "r"(c9), "r"(c10));
#else
    {
        uint64_t s = (uint64_t)o[7] + e[8] + c8;
        x[7] = (uint32_t)s;
        s = (uint64_t)e[9] + c9 + (s >> 32);
        t9 = (uint32_t)s;
        t10 = c10 + (uint32_t)(s >> 32);
    }
#endif
    redc11<F>(x, y, t9, t10);      // row 1: y[0] -> 0
    slide11(y, x[0], t9, t10);     // y becomes hi-aligned, x lo-aligned
    redc11<F>(y, x, t9, t10);      // row 2: x[0] -> 0
    slide11(x, y[0], t9, t10);     // x hi-aligned, y lo-aligned
    redc11<F>(x, y, t9, t10);      // row 3: y[0] -> 0
    // value = x + (y >> 32) < 2p < 2^256 (columns 9 and 10 are empty again): one conditional subtraction
    uint32_t s8[8];
    merge8(s8, x, y);
    cond_sub_p<F>(r, s8);
}


// r = sum_{j<N} a[j] * c_j  (Montgomery in -> Montgomery out); `entry(j, dgt, t)` fills t[0..7] with
// entry dgt of the table of c_j (read where it is used, so only one entry is live at a time)
template <class F, int N, class Entry>
PSD_DEV void const_dot(uint32_t* r, const uint32_t (*a)[8], Entry entry) {
    uint32_t e[10], o[8], c8 = 0, c9 = 0, c10 = 0;
#pragma unroll
    for (int k = 0; k < 10; k++) e[k] = 0;
#pragma unroll
    for (int k = 0; k < 8; k++) o[k] = 0;
#pragma unroll
    for (int j = 0; j < N; j++)
#pragma unroll
        for (int d = 0; d < 4; d++) {
            uint32_t t[8];
            entry(j, d, t);
            cacc_digit(e, o, c8, c9, c10, a[j][2 * d], a[j][2 * d + 1], t);
        }
    cacc_finish<F>(r, e, o, c8, c9, c10);
}

// b = min(a, p - a) for a in [0, p)  (b*b == a*a mod p)
template <class F>
PSD_DEV void neg_if_larger_half(uint32_t* b, const uint32_t* a) {
    uint32_t n[8], borrow;
#if !defined(PSD_HOST_EMU)
    // n = p - a ; borrow of (n - a) tells whether n < a, i.e. a > p/2
    asm("{\n\t"
        ".reg .u32 d;\n\t"
        "sub.cc.u32 %0, %9, %17;\n\t"
        "subc.cc.u32 %1, %10, %18;\n\t"
        "subc.cc.u32 %2, %11, %19;\n\t"
        "subc.cc.u32 %3, %12, %20;\n\t"
        "subc.cc.u32 %4, %13, %21;\n\t"
        "subc.cc.u32 %5, %14, %22;\n\t"
        "subc.cc.u32 %6, %15, %23;\n\t"
        "subc.u32 %7, %16, %24;\n\t"
        "sub.cc.u32 d, %0, %17;\n\t"
        "subc.cc.u32 d, %1, %18;\n\t"
        "subc.cc.u32 d, %2, %19;\n\t"
        "subc.cc.u32 d, %3, %20;\n\t"
        "subc.cc.u32 d, %4, %21;\n\t"
        "subc.cc.u32 d, %5, %22;\n\t"
        "subc.cc.u32 d, %6, %23;\n\t"
        "subc.cc.u32 d, %7, %24;\n\t"
        "subc.u32 %8, 0, 0;\n\t"
        "}"
        : "=&r"(n[0]), "=&r"(n[1]), "=&r"(n[2]), "=&r"(n[3]), "=&r"(n[4]), "=&r"(n[5]), "=&r"(n[6]), "=&r"(n[7]),
          "=&r"(borrow)
        : "r"(plimb<F, 0>()), "r"(plimb<F, 1>()), "r"(plimb<F, 2>()), "r"(plimb<F, 3>()), "r"(plimb<F, 4>()),
          "r"(plimb<F, 5>()), "r"(plimb<F, 6>()), "r"(plimb<F, 7>()),
          "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(a[4]), "r"(a[5]), "r"(a[6]), "r"(a[7]));
#else
    const uint32_t p[8] = {F::P0, F::P1, F::P2, F::P3, F::P4, F::P5, F::P6, F::P7};
    int64_t c = 0;
    for (int k = 0; k < 8; k++) {
        int64_t t = (int64_t)p[k] - a[k] + c;
        n[k] = (uint32_t)t;
        c = t >> 32;
    }
    c = 0;
    for (int k = 0; k < 8; k++) {
        int64_t t = (int64_t)n[k] - a[k] + c;
        c = t >> 32;
    }
    borrow = (uint32_t)c;
#endif
    // borrow != 0  <=>  p - a < a  <=>  take n
#pragma unroll
    for (int k = 0; k < 8; k++) b[k] = borrow ? n[k] : a[k];
}

// Squaring entry point used by the S-box: dedicated squaring where its bound holds.
template <class F, bool REDUCE = true>
PSD_DEV void fp_sqr(uint32_t* r, const uint32_t* a) {
    if (F::kBits == 254) {
        mont_sqr<F, REDUCE>(r, a);
    } else {
        // BLS12-381: 3p > 2^256, so mont_sqr's 9-limb window could overflow for a close to p; the wide
        // variant (mont_sqr_w) works but its carry handling makes it slower than a plain product on
        // B200 (6.49e10 squarings/s vs 7.06e10 products/s).  Instead square the smaller of a and p - a
        // (same square mod p): b <= (p-1)/2 gives 2b + p < 2p < 2^256 and the narrow squaring applies.
        uint32_t b[8];
        neg_if_larger_half<F>(b, a);
        mont_sqr<F>(r, b);
    }
}

// x^alpha for the S-box (apply_s_box, poseidon/mod.rs:63-74).  ark-ff's `pow` is an MSB-first
// square-and-multiply; exponentiation is exponentiation, so any chain gives the same residue.
template <class F>
PSD_DEV void fp_pow5(uint32_t* r, const uint32_t* x) {
    uint32_t x2[8], x4[8];
    fp_sqr<F>(x2, x);
    fp_sqr<F>(x4, x2);
    mont_mul<F>(r, x4, x);
}

template <class F>
PSD_DEV void fp_pow_u64(uint32_t* r, const uint32_t* x, uint64_t alpha, const uint32_t* one_mont) {
    // generic MSB-first ladder; alpha is uniform across the grid
    uint32_t acc[8];
#pragma unroll
    for (int k = 0; k < 8; k++) acc[k] = one_mont[k];
    bool started = false;
#pragma unroll 1
    for (int bit = 63; bit >= 0; bit--) {
        if (started) fp_sqr<F>(acc, acc);
        if ((alpha >> bit) & 1) {
            if (started) {
                mont_mul<F>(acc, acc, x);
            } else {
#pragma unroll
                for (int k = 0; k < 8; k++) acc[k] = x[k];
                started = true;
            }
        }
    }
#pragma unroll
    for (int k = 0; k < 8; k++) r[k] = acc[k];
}

}  // namespace psd
