// Host-side (one-time) tables for the FP64-pipe linear layer (fp52.cuh, poseidon_fp52.cuh).
//
// Everything here is derived from the regrouped constants of opt_constants.cpp (M_pre, the sparse
// rows (M00, w), the lane coefficients v, the per-round constants k_i and post_ark).  A constant c
// (true value) that multiplies a lane becomes the table  T_l = c * 2^(e + 52 l) mod p, l = 0..4, each
// entry split into five 52-bit limbs stored as doubles; e collects the Montgomery rows that follow
// (2 for a row, 0 for a lane update) and the lane scale sigma_i = 2^(-52 i) (see poseidon_fp52.cuh).
// Additive constants are the six starting values of the column sums: the constant's residue times
// 2^104, minus the IEEE exponent patterns that the products will add (fp52.cuh).
#include "fp52_tables.hpp"

#include <string.h>

namespace psd {

namespace {

constexpr uint64_t kMask52h = (1ull << 52) - 1;
constexpr uint64_t kLo = 0x4330000000000000ull, kHi = 0x4670000000000000ull;

struct Pow2 {
    const HostField& f;
    Fe two, half;
    explicit Pow2(const HostField& hf) : f(hf) {
        Fe t = {{2, 0, 0, 0}};
        two = hf_to_mont(f, t);
        half = hf_inv(f, two);
    }
    // Montgomery form of 2^e mod p, e may be negative
    Fe operator()(long e) const { return e >= 0 ? hf_pow_u64(f, two, (uint64_t)e) : hf_pow_u64(f, half, (uint64_t)(-e)); }
};

void split52(const Fe& x, uint64_t out[5]) {
    // bits 52 l .. 52 l + 51 of the 256-bit value
    for (int l = 0; l < 5; l++) {
        const int bit = 52 * l, w = bit / 64, s = bit % 64;
        uint64_t v = x.v[w] >> s;
        if (s > 12 && w + 1 < 4) v |= x.v[w + 1] << (64 - s);
        out[l] = v & kMask52h;
    }
}

void push_u64(std::vector<uint32_t>& img, uint64_t q) {
    img.push_back((uint32_t)q);
    img.push_back((uint32_t)(q >> 32));
}
void push_double(std::vector<uint32_t>& img, double d) {
    uint64_t q;
    memcpy(&q, &d, 8);
    push_u64(img, q);
}

// table of the constant with Montgomery form c_mont, exponent base e
void push_table(std::vector<uint32_t>& img, const HostField& f, const Pow2& pow2, const Fe& c_mont, long e) {
    for (int l = 0; l < 5; l++) {
        // (c R)(2^x R)/R = (c 2^x) R  -> canonical
        const Fe canon = hf_from_mont(f, hf_mul(f, c_mont, pow2(e + 52 * l)));
        uint64_t limbs[5];
        split52(canon, limbs);
        for (int k = 0; k < 5; k++) push_double(img, (double)limbs[k]);
    }
}

// six column start values for a row over t lanes whose additive constant has the residue k_res
void push_k(std::vector<uint32_t>& img, const HostField& f, const Pow2& pow2, const Fe& k_res, uint32_t t) {
    // residue (as an integer) * 2^104 mod p:  a * (2^104 R) / R
    const Fe v = hf_mul(f, k_res, pow2(104));
    uint64_t limbs[5];
    split52(v, limbs);
    // offsets each start column will receive: 5t products, then two Montgomery rows with a column
    // shift after each (fp52.cuh f52_mont_row)
    const uint64_t n = 5ull * t;
    push_u64(img, limbs[0] - (n + 1) * kLo);
    push_u64(img, limbs[1] - ((n + 2) * kLo + (n + 1) * kHi));
    for (int l = 2; l < 5; l++) push_u64(img, limbs[l] - (n + 2) * (kLo + kHi));
    push_u64(img, 0 - ((n + 2) * kHi + kLo));
}

Fe fe_at(const std::vector<uint32_t>& img, uint32_t word_off) {
    Fe x;
    for (int k = 0; k < 4; k++) x.v[k] = (uint64_t)img[word_off + 2 * k] | ((uint64_t)img[word_off + 2 * k + 1] << 32);
    return x;
}

}  // namespace

void append_fp52_tables(const HostField& f, uint32_t t, uint32_t full_rounds, uint32_t partial_rounds,
                        const std::vector<Fe>& ark, const std::vector<Fe>& mds, PermDesc& d,
                        std::vector<uint32_t>& img) {
    d.fp_ok = 0;
    d.fp_consts_off = d.fp_mat_off = d.fp_pre_off = d.fp_last_off = d.fp_k_off = d.fp_part_off = 0;
    const uint32_t half = full_rounds / 2, rp = partial_rounds, n = t - 1, total = full_rounds + rp;
    if (half == 0 || full_rounds - half == 0) return;

    // the regrouped constants, from a scratch image
    PermDesc od{};
    std::vector<uint32_t> oimg;
    append_optimised_constants(f, t, full_rounds, rp, ark, mds, od, oimg);
    if (!od.opt_ok) return;
    std::vector<Fe> m_pre, k_lane0(rp), post_ark(t);
    std::vector<std::vector<Fe>> w(rp), v(rp);
    if (rp) {
        for (uint32_t i = 0; i < t * t; i++) m_pre.push_back(fe_at(oimg, od.pre_mds_off + i * 8));
        for (uint32_t r = 0; r < rp; r++) {
            const uint32_t base = od.sparse_off + r * 2 * t * 8;
            k_lane0[r] = fe_at(oimg, base);
            for (uint32_t j = 0; j < n; j++) w[r].push_back(fe_at(oimg, base + (2 + j) * 8));
            for (uint32_t j = 0; j < n; j++) v[r].push_back(fe_at(oimg, base + (2 + n + j) * 8));
        }
        for (uint32_t j = 0; j < t; j++) post_ark[j] = fe_at(oimg, od.post_ark_off + j * 8);
    }

    const Pow2 pow2(f);
    while (img.size() % 8) img.push_back(0);
    auto here = [&img]() { return (uint32_t)(img.size() / 2); };

    // consts: P52[5], NP52, SHIFT, 2^104
    d.fp_consts_off = here();
    {
        Fe p;
        memcpy(p.v, f.p, 32);
        uint64_t limbs[5];
        split52(p, limbs);
        for (int k = 0; k < 5; k++) push_double(img, (double)limbs[k]);
        push_u64(img, f.np0 & kMask52h);
        push_u64(img, 0 - kHi);
        push_u64(img, kHi);   // the double 2^104
    }
    // matrices
    d.fp_mat_off = here();
    for (uint32_t i = 0; i < t * t; i++) push_table(img, f, pow2, mds[i], 104);
    if (rp) {
        d.fp_pre_off = here();
        for (uint32_t i = 0; i < t * t; i++) push_table(img, f, pow2, m_pre[i], 104);
        // last partial round, lane scale sigma = 2^(-52 (rp-1)):
        //   row 0: M00, w_j / sigma;   row j: v_j, then 1/sigma on lane j and 0 elsewhere
        d.fp_last_off = here();
        const long up = 52L * (rp - 1);
        const Fe one = hf_one(f), zero = hf_zero();
        push_table(img, f, pow2, mds[0], 104);
        for (uint32_t j = 0; j < n; j++) push_table(img, f, pow2, w[rp - 1][j], 104 + up);
        for (uint32_t r = 0; r < n; r++) {
            push_table(img, f, pow2, v[rp - 1][r], 104);
            for (uint32_t j = 0; j < n; j++) push_table(img, f, pow2, j == r ? one : zero, 104 + up);
        }
    }
    // additive constants in execution order: what round r's rows add is what round r+1 starts with
    d.fp_k_off = here();
    const Fe zero = hf_zero();
    for (uint32_t r = 0; r < total; r++) {
        const bool full = r < half || r >= half + rp;
        const bool last_partial = !full && r + 1 == half + rp;
        if (full || last_partial) {
            for (uint32_t j = 0; j < t; j++) {
                Fe k = zero;
                if (last_partial) k = post_ark[j];
                else if (r + 1 == half && rp) k = (j == 0) ? k_lane0[0] : zero;
                else if (r + 1 < total) k = ark[(size_t)(r + 1) * t + j];
                push_k(img, f, pow2, k, t);
            }
        } else {
            push_k(img, f, pow2, k_lane0[r - half + 1], t);
        }
    }
    // every partial round: w tables (row 0, two Montgomery rows) and v tables (lane update; the last
    // round's are never read -- its lanes go through the `last` matrix -- but keep the stride)
    d.fp_part_off = here();
    for (uint32_t i = 0; i < rp; i++) {
        for (uint32_t j = 0; j < n; j++) push_table(img, f, pow2, w[i][j], 104 + 52L * i);
        for (uint32_t j = 0; j < n; j++) push_table(img, f, pow2, v[i][j], -52L * i);
    }
    while (img.size() % 8) img.push_back(0);
    d.fp_ok = 1;
    // the kernel's exactness rests on bounds that depend on these very numbers: evaluate them now
    if (!fp52_image_bounds_ok(f, t, full_rounds, partial_rounds, d, img, nullptr)) d.fp_ok = 0;
}

// ------------------------------------------------------------------------------------------------
// Bounds behind variant 5, evaluated for the image that is about to be uploaded.
//
// What the device code assumes (fp52.cuh, poseidon_fp52.cuh):
//  (1) every table / modulus limb is an integer in [0, 2^52) -- otherwise the two-FMA product is not exact --
//      and every 5-limb table entry is a canonical residue (< p);
//  (2) the additive constants, with the IEEE exponent patterns added back, are limbs < 2^52 of a value < p;
//  (3) a row over lanes with the limb bounds below returns less than p + 2^207, and that is a valid S-box input
//      (< 2^256, inside the lazy S-box's operand range);
//  (4) the lazily reduced lanes of the partial rounds stay below 2^258 (top limb < 2^50) for every round's
//      lane-update tables;
//  (5) no 64-bit column sum wraps;
//  (6) the lazy S-box maps [0, p + 2^207) into values below 2^256 whose squarings respect 2a + p < 2^256.
// The sums are taken with the worst case of every limb (2^52 - 1, or the top-limb bound) against the ACTUAL
// table entries, in long double with every product rounded up by 2^-40 -- the slack of each inequality is
// at least 2^-6 relative, so the floating-point evaluation decides them safely.
// ------------------------------------------------------------------------------------------------
namespace {

const long double kUp = 1.0L + 0x1p-40L;

struct ImageReader {
    const std::vector<uint32_t>& img;
    bool in_range(uint32_t off8, uint32_t n) const { return ((size_t)off8 + n) * 2 <= img.size(); }
    uint64_t u64(uint32_t off8) const { return (uint64_t)img[2 * (size_t)off8] | ((uint64_t)img[2 * (size_t)off8 + 1] << 32); }
    double dbl(uint32_t off8) const {
        const uint64_t q = u64(off8);
        double x;
        memcpy(&x, &q, 8);
        return x;
    }
};

// value of a 5-limb table entry, as a long double for the sums and exactly (256 bits) for comparisons with p;
// false if a limb is not an integer in [0, 2^52) or the value needs more than 256 bits
bool entry_value(const ImageReader& r, uint32_t off8, long double* val, Fe* exact = nullptr) {
    long double v = 0;
    uint64_t limbs[5];
    for (int k = 4; k >= 0; k--) {
        const double x = r.dbl(off8 + k);
        if (!(x >= 0.0) || x >= 4503599627370496.0 || x != (double)(uint64_t)x) return false;
        limbs[k] = (uint64_t)x;
        v = v * 0x1p52L + (long double)x;
    }
    if (limbs[4] >> 48) return false;
    if (exact) {
        exact->v[0] = limbs[0] | (limbs[1] << 52);
        exact->v[1] = (limbs[1] >> 12) | (limbs[2] << 40);
        exact->v[2] = (limbs[2] >> 24) | (limbs[3] << 28);
        exact->v[3] = (limbs[3] >> 36) | (limbs[4] << 16);
    }
    *val = v;
    return true;
}

// worst case of sum_i X_i * entry_i over one table (5 entries), X_i <= lim[i]
bool table_worst(const ImageReader& r, uint32_t tab_off8, const long double lim[5], const HostField& f, long double* sum) {
    long double s = 0;
    for (int i = 0; i < 5; i++) {
        long double v;
        Fe exact;
        if (!entry_value(r, tab_off8 + 5 * i, &v, &exact)) return false;
        if (hf_cmp(exact.v, f.p) >= 0) return false;   // not a residue
        s += lim[i] * v * kUp;
    }
    *sum = s;
    return true;
}

}  // namespace

bool fp52_image_bounds_ok(const HostField& f, uint32_t t, uint32_t full_rounds, uint32_t partial_rounds, const PermDesc& d,
                          const std::vector<uint32_t>& img, std::string* why) {
    auto no = [&](const char* msg) {
        if (why) *why = msg;
        return false;
    };
    const ImageReader r{img};
    const uint32_t rp = partial_rounds, n = t - 1, half = full_rounds / 2;
    const long double two52 = 0x1p52L, two104 = 0x1p104L, lim52 = two52 - 1;
    long double p = 0;
    for (int k = 3; k >= 0; k--) p = p * 0x1p64L + (long double)f.p[k];
    const bool lazy = f.bits == 254;
    if (t < 2 || t > 5) return no("width outside [2,5]");
    if ((10.0L * t + 8) * two52 >= 0x1p63L) return no("column sums could wrap");                       // (5)

    // consts: P52[5], NP52, SHIFT, 2^104
    if (!r.in_range(d.fp_consts_off, 8)) return no("constants outside the image");
    {
        long double pv;
        Fe pe;
        if (!entry_value(r, d.fp_consts_off, &pv, &pe) || hf_cmp(pe.v, f.p) != 0) return no("P52 is not the modulus");   // (1)
        const uint64_t np = r.u64(d.fp_consts_off + 5), mask = (1ull << 52) - 1;
        if (np > mask || ((np * (f.p[0] & mask) + 1) & mask) != 0) return no("NP52 * p != -1 mod 2^52");
        if (r.u64(d.fp_consts_off + 6) != 0 - kHi || r.u64(d.fp_consts_off + 7) != kHi) return no("SHIFT / 2^104 constants");
    }

    // limb bounds of the values a row can meet
    //   S-box output: < 2^256 (lazy: < 1.26 p)              -> limbs 0..3 <= 2^52-1, limb 4 < 2^48
    //   row output / fresh limb state (unpacked 256-bit)    -> the same
    //   lazily reduced lane of the partial rounds: < B_i    -> limb 4 <= B_i / 2^208
    long double lim_full[5] = {lim52, lim52, lim52, lim52, 0x1p48L - 1};
    const long double row_limit = p + 0x1p207L;                                                          // (3)
    if (row_limit >= 0x1p256L) return no("p + 2^207 does not fit 256 bits");
    auto row_ok = [&](uint32_t off_first, uint32_t off_rest, const long double (*lims)[5], long double* worst) {
        long double s = p;   // additive constant: a residue < p
        for (uint32_t j = 0; j < t; j++) {
            const uint32_t off = j == 0 ? off_first : off_rest + (j - 1) * 25;
            if (!r.in_range(off, 25)) return false;
            long double part;
            if (!table_worst(r, off, lims[j], f, &part)) return false;
            s += part;
        }
        // The two Montgomery rows add m1 p + m2 p 2^52 with m1, m2 < 2^52: the second contributes less than p
        // to the result, so (result - p) < (constant + products + m1 p) / 2^104, and THAT is what has to stay
        // below 2^207 -- a comparison with a factor of slack, unlike result < p + 2^207 itself, whose two
        // sides agree to 46 bits.
        s += lim52 * p * kUp;
        *worst = s / two104 * kUp;
        return *worst < 0x1p207L;
    };
    long double lims_full[5][5];
    for (uint32_t j = 0; j < 5; j++) memcpy(lims_full[j], lim_full, sizeof lim_full);
    long double worst = 0;
    // dense matrices: mat always, pre / last when there are partial rounds
    for (uint32_t i = 0; i < t; i++) {
        const uint32_t off = d.fp_mat_off + i * t * 25;
        if (!row_ok(off, off + 25, lims_full, &worst)) return no("a row of the MDS tables can exceed p + 2^207");
    }
    // additive constants (2): [rounds] x 6 start values; undo the offsets push_k subtracted
    {
        uint32_t n_k = 0;
        const uint32_t total = full_rounds + rp;
        for (uint32_t rr = 0; rr < total; rr++) {
            const bool full = rr < half || rr >= half + rp;
            const bool last_partial = !full && rr + 1 == half + rp;
            n_k += (full || last_partial) ? t : 1;
        }
        if (!r.in_range(d.fp_k_off, n_k * 6)) return no("additive constants outside the image");
        const uint64_t nn = 5ull * t;
        for (uint32_t e = 0; e < n_k; e++) {
            uint64_t limbs[5];
            limbs[0] = r.u64(d.fp_k_off + 6 * e) + (nn + 1) * kLo;
            limbs[1] = r.u64(d.fp_k_off + 6 * e + 1) + ((nn + 2) * kLo + (nn + 1) * kHi);
            for (int l = 2; l < 5; l++) limbs[l] = r.u64(d.fp_k_off + 6 * e + l) + (nn + 2) * (kLo + kHi);
            if (r.u64(d.fp_k_off + 6 * e + 5) != 0 - ((nn + 2) * kHi + kLo)) return no("top column start value");
            for (int l = 4; l >= 0; l--)
                if (limbs[l] > kMask52h) return no("an additive constant's limb is not below 2^52");
            if (limbs[4] >> 48) return no("an additive constant is not a residue");
            Fe kv;
            kv.v[0] = limbs[0] | (limbs[1] << 52);
            kv.v[1] = (limbs[1] >> 12) | (limbs[2] << 40);
            kv.v[2] = (limbs[2] >> 24) | (limbs[3] << 28);
            kv.v[3] = (limbs[3] >> 36) | (limbs[4] << 16);
            if (hf_cmp(kv.v, f.p) >= 0) return no("an additive constant is not a residue");
        }
    }
    if (rp) {
        for (uint32_t i = 0; i < t; i++) {
            const uint32_t off = d.fp_pre_off + i * t * 25;
            if (!row_ok(off, off + 25, lims_full, &worst)) return no("a row of the M_pre tables can exceed p + 2^207");
        }
        // partial rounds: lanes j >= 1 start as row outputs (< p + 2^207) and follow B' = (B + sum + m p) / 2^52  (4)
        std::vector<long double> B(n, row_limit);
        long double lim_sbox[5];
        memcpy(lim_sbox, lim_full, sizeof lim_sbox);
        for (uint32_t i = 0; i < rp; i++) {
            const uint32_t part = d.fp_part_off + i * 2 * n * 25;
            if (!r.in_range(part, 2 * n * 25)) return no("partial-round tables outside the image");
            long double lims[5][5];
            memcpy(lims[0], lim_sbox, sizeof lim_sbox);
            for (uint32_t j = 0; j < n; j++) {
                memcpy(lims[j + 1], lim_full, sizeof lim_full);
                lims[j + 1][4] = B[j] / 0x1p208L * kUp;
                if (lims[j + 1][4] >= 0x1p50L) return no("a lazily reduced lane can reach 2^258");
            }
            // row 0 of this round: M00 table (mat[0][0]) + this round's w tables
            if (!row_ok(d.fp_mat_off, part, lims, &worst)) return no("row 0 of a partial round can exceed p + 2^207");
            if (i + 1 < rp) {
                // lane updates with this round's v tables
                for (uint32_t j = 0; j < n; j++) {
                    long double s;
                    if (!table_worst(r, part + (n + j) * 25, lim_sbox, f, &s)) return no("a lane-update table entry is out of range");
                    B[j] = (B[j] + s + lim52 * p * kUp) / two52 * kUp;
                    if (B[j] >= 0x1p258L) return no("a lazily reduced lane can reach 2^258");
                }
            } else {
                // last partial round: rows 1..t-1 come from the `last` matrix over (S-box output, lazy lanes)
                for (uint32_t row = 0; row < t; row++) {
                    const uint32_t off = d.fp_last_off + row * t * 25;
                    if (!row_ok(off, off + 25, lims, &worst)) return no("a row of the last partial round can exceed p + 2^207");
                }
            }
        }
    }
    // (6) the S-box on [0, p + 2^207): lazy (BN254) -- no reductions inside x^alpha; squaring operands need
    // 2a + p < 2^256, the closing product a + p < 2^256, and the result must fit 256 bits
    if (lazy) {
        const long double q = p / 0x1p256L;
        const long double x0 = row_limit / p;
        long double y = x0;   // in units of p
        for (int sq = 0; sq < 4; sq++) {   // alpha = 17 has the most squarings; alpha = 5 stops after two
            if ((2 * y + 1) * p >= 0x1p256L) return no("a squaring operand leaves the 9-limb window");
            y = (1 + y * y * q) * kUp;
        }
        if ((y + 1) * p >= 0x1p256L) return no("the closing product's operand is too large");
        const long double out = (1 + y * x0 * q) * kUp;
        if (out * p >= 0x1p256L || out > 1.26L) return no("the lazy S-box can exceed 1.26 p");
    } else if (2 * p <= row_limit) {
        return no("row outputs are not below 2p");
    }
    return true;
}

}  // namespace psd
