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
}

}  // namespace psd
