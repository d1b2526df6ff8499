// Host build of sponge_b200/csrc/fp256.cuh with the C carry-chain model (PSD_HOST_EMU), exported
// for ctypes so the limb plumbing of the CUDA arithmetic can be checked without a GPU.
// Test infrastructure only.
#define PSD_HOST_EMU 1
#include "../../sponge_b200/csrc/fp256.cuh"
#include <stddef.h>
#include <string.h>

using namespace psd;

extern "C" {
void emu_mont_mul(int field, uint32_t* r, const uint32_t* a, const uint32_t* b, size_t n) {
    for (size_t i = 0; i < n; i++) {
        if (field == 0) mont_mul<Bn254Fr>(r + 8 * i, a + 8 * i, b + 8 * i);
        else mont_mul<Bls12381Fr>(r + 8 * i, a + 8 * i, b + 8 * i);
    }
}
void emu_add(int field, uint32_t* r, const uint32_t* a, const uint32_t* b, size_t n) {
    for (size_t i = 0; i < n; i++) {
        if (field == 0) fp_add<Bn254Fr>(r + 8 * i, a + 8 * i, b + 8 * i);
        else fp_add<Bls12381Fr>(r + 8 * i, a + 8 * i, b + 8 * i);
    }
}
void emu_pow(int field, uint32_t* r, const uint32_t* a, uint64_t alpha, const uint32_t* one, size_t n) {
    for (size_t i = 0; i < n; i++) {
        if (field == 0) fp_pow_u64<Bn254Fr>(r + 8 * i, a + 8 * i, alpha, one);
        else fp_pow_u64<Bls12381Fr>(r + 8 * i, a + 8 * i, alpha, one);
    }
}
}

extern "C" void emu_mont_dot(int n_terms, uint32_t* r, const uint32_t* a, const uint32_t* b, size_t n) {
    // BN254 only (the no-overflow bound needs (N+1)p < 2^256); a, b: n * n_terms * 8 words
    for (size_t i = 0; i < n; i++) {
        const uint32_t(*ap)[8] = reinterpret_cast<const uint32_t(*)[8]>(a + i * n_terms * 8);
        const uint32_t(*bp)[8] = reinterpret_cast<const uint32_t(*)[8]>(b + i * n_terms * 8);
        switch (n_terms) {
            case 1: mont_dot<Bn254Fr, 1>(r + 8 * i, ap, bp); break;
            case 2: mont_dot<Bn254Fr, 2>(r + 8 * i, ap, bp); break;
            case 3: mont_dot<Bn254Fr, 3>(r + 8 * i, ap, bp); break;
            case 4: mont_dot<Bn254Fr, 4>(r + 8 * i, ap, bp); break;
        }
    }
}

// ---- whole permutation on the host model (dense vs optimised), constants from the product's own
// host code (params.cpp / opt_constants.cpp) ------------------------------------------------------
#include <vector>

#include <fenv.h>

#include "../../sponge_b200/csrc/fp52_tables.hpp"
#include "../../sponge_b200/csrc/opt_constants.hpp"
#include "../../sponge_b200/csrc/poseidon_core.cuh"
#include "../../sponge_b200/csrc/poseidon_fp52.cuh"

// the FP64-pipe variant exists for alpha 5 / 17 only
template <class F, int T, int ALPHA>
static bool emu_perm_fp52(uint32_t (&s)[T][8], const PermDesc& d) {
    if constexpr ((ALPHA == 5 || ALPHA == 17) && T <= 5) {
        permute_fp52<F, T, ALPHA>(s, d);
        return true;
    } else {
        return false;
    }
}

template <class F, int T, int ALPHA>
static void emu_perm_t(const PermDesc& d, int variant, uint32_t* st, size_t n) {
    for (size_t i = 0; i < n; i++) {
        uint32_t s[T][8];
        memcpy(s, st + i * T * 8, sizeof s);
        if (variant == 1 && d.opt_ok) permute_opt<F, T, ALPHA>(s, d);
        else if (variant == 2) permute_rolled<F, T, ALPHA>(s, d);
        else if (variant == 3 && d.opt_ok) permute_opt_compact<F, T, ALPHA>(s, d);
        else if (variant == 4 && d.ctab_ok == 1) permute_opt_ctab<F, T, ALPHA>(s, d);
        else if (variant == 5 && d.ctab_ok == 2) permute_opt_ctab<F, T, ALPHA, true>(s, d);
        else if (variant == 6 && d.fp_ok && emu_perm_fp52<F, T, ALPHA>(s, d)) {}
        else permute_dense<F, T, ALPHA>(s, d);
        memcpy(st + i * T * 8, s, sizeof s);
    }
}
template <class F, int T>
static void emu_perm_a(const PermDesc& d, uint64_t alpha, int variant, uint32_t* st, size_t n) {
    if (alpha == 5) emu_perm_t<F, T, 5>(d, variant, st, n);
    else if (alpha == 17) emu_perm_t<F, T, 17>(d, variant, st, n);
    else emu_perm_t<F, T, 0>(d, variant, st, n);
}
template <class F>
static int emu_perm_f(const PermDesc& d, uint32_t t, uint64_t alpha, int variant, uint32_t* st, size_t n) {
    switch (t) {
        case 2: emu_perm_a<F, 2>(d, alpha, variant, st, n); return 0;
        case 3: emu_perm_a<F, 3>(d, alpha, variant, st, n); return 0;
        case 4: emu_perm_a<F, 4>(d, alpha, variant, st, n); return 0;
        case 5: emu_perm_a<F, 5>(d, alpha, variant, st, n); return 0;
        case 9: emu_perm_a<F, 9>(d, alpha, variant, st, n); return 0;
    }
    return -1;
}

// returns 0 ok, -1 unsupported t, 1 if the optimised form was requested but is not available
extern "C" int emu_permute(int field, uint32_t rate, uint32_t capacity, uint32_t rf, uint32_t rp, uint64_t alpha,
                           const uint64_t* ark_canon, const uint64_t* mds_canon, int variant, uint32_t* states,
                           size_t n) {
    const HostField* hf = host_field(field);
    const uint32_t t = rate + capacity, rounds = rf + rp;
    std::vector<Fe> ark(rounds * t), mds(t * t);
    for (size_t i = 0; i < ark.size(); i++) ark[i] = hf_to_mont(*hf, reinterpret_cast<const Fe*>(ark_canon)[i]);
    for (size_t i = 0; i < mds.size(); i++) mds[i] = hf_to_mont(*hf, reinterpret_cast<const Fe*>(mds_canon)[i]);
    std::vector<uint32_t> img;
    auto push = [&img](const Fe& f) {
        for (int k = 0; k < 4; k++) { img.push_back((uint32_t)f.v[k]); img.push_back((uint32_t)(f.v[k] >> 32)); }
    };
    PermDesc d{};
    d.ark_off = 0;
    // variant 6 only reads the round-0 constants from here (as the library's variant-5 image does)
    for (size_t i = 0; i < (variant == 6 ? (size_t)t : ark.size()); i++) push(ark[i]);
    d.mds_off = (uint32_t)img.size();
    for (auto& f : mds) push(f);
    d.one_off = (uint32_t)img.size();
    push(hf_one(*hf));
    d.r2_off = (uint32_t)img.size();
    Fe r2; memcpy(r2.v, hf->r2, 32); push(r2);
    d.half_full = rf / 2; d.partial = rp; d.total = rounds; d.rate = rate; d.capacity = capacity;
    d.alpha_lo = (uint32_t)alpha; d.alpha_hi = (uint32_t)(alpha >> 32);
    if (variant != 6) append_optimised_constants(*hf, t, rf, rp, ark, mds, d, img);
    if (variant == 4) append_digit_tables(*hf, t, rf, rp, mds, d, img);
    if (variant == 5) append_digit_tables(*hf, t, rf, rp, mds, d, img, true);
    if (variant == 6) append_fp52_tables(*hf, t, rf, rp, ark, mds, d, img);
    if (img.size() > kSmemPoolMaxWords) return variant == 6 ? 1 : -2;   // the host model's pool has the shared-memory form's size
    memcpy(c_pool, img.data(), img.size() * 4);
    // fp52.cuh's host model is fma() under round-toward-zero (the device uses fma.rz.f64)
    const int old_round = fegetround();
    fesetround(FE_TOWARDZERO);
    int rc = field == 0 ? emu_perm_f<Bn254Fr>(d, t, alpha, variant, states, n)
                        : emu_perm_f<Bls12381Fr>(d, t, alpha, variant, states, n);
    fesetround(old_round);
    if (rc) return rc;
    if (variant == 6 && !(d.fp_ok && (alpha == 5 || alpha == 17) && t <= 5)) return 1;
    if ((variant == 4 && d.ctab_ok != 1) || (variant == 5 && d.ctab_ok != 2)) return 1;
    return ((variant == 1 || variant == 3) && !d.opt_ok) ? 1 : 0;
}

// BN254 S-box without its inner conditional subtractions (sbox_compact's LAZY mode): raw 256-bit result
extern "C" void emu_sbox_lazy_bn254(uint64_t alpha, uint32_t* r, const uint32_t* a, size_t n) {
    PermDesc d{};
    for (size_t i = 0; i < n; i++) {
        uint32_t x[8];
        memcpy(x, a + 8 * i, 32);
        if (alpha == 5) sbox_compact<Bn254Fr, 5, true>(x, d);
        else sbox_compact<Bn254Fr, 17, true>(x, d);
        memcpy(r + 8 * i, x, 32);
    }
}

extern "C" void emu_mont_sqr(int field, uint32_t* r, const uint32_t* a, size_t n) {
    for (size_t i = 0; i < n; i++) {
        if (field == 0) mont_sqr<Bn254Fr>(r + 8 * i, a + 8 * i);
        else fp_sqr<Bls12381Fr>(r + 8 * i, a + 8 * i);
    }
}

extern "C" void emu_mont_dot_w(int field, int n_terms, uint32_t* r, const uint32_t* a, const uint32_t* b, size_t n) {
    for (size_t i = 0; i < n; i++) {
        const uint32_t(*ap)[8] = reinterpret_cast<const uint32_t(*)[8]>(a + i * n_terms * 8);
        const uint32_t(*bp)[8] = reinterpret_cast<const uint32_t(*)[8]>(b + i * n_terms * 8);
#define PSD_CASE(N)                                                             \
    case N:                                                                     \
        if (field == 0) mont_dot_w<Bn254Fr, N>(r + 8 * i, ap, bp);              \
        else mont_dot_w<Bls12381Fr, N>(r + 8 * i, ap, bp);                      \
        break;
        switch (n_terms) { PSD_CASE(1) PSD_CASE(2) PSD_CASE(3) PSD_CASE(4) }
#undef PSD_CASE
    }
}
extern "C" void emu_mont_sqr_w(int field, uint32_t* r, const uint32_t* a, size_t n) {
    for (size_t i = 0; i < n; i++) {
        if (field == 0) mont_sqr_w<Bn254Fr>(r + 8 * i, a + 8 * i);
        else mont_sqr_w<Bls12381Fr>(r + 8 * i, a + 8 * i);
    }
}

extern "C" void emu_mont_dot_w5_bn254(uint32_t* r, const uint32_t* a, const uint32_t* b, size_t n) {
    for (size_t i = 0; i < n; i++)
        mont_dot_w<Bn254Fr, 5>(r + 8 * i, reinterpret_cast<const uint32_t(*)[8]>(a + i * 40),
                               reinterpret_cast<const uint32_t(*)[8]>(b + i * 40));
}

// digit-table constant products: tables built here with the host field (tab[d] = c * 2^(64d+96) mod p)
static void build_digit_table(const HostField& hf, const uint32_t* c_mont, uint32_t (*tab)[8]) {
    Fe c;
    memcpy(c.v, c_mont, 32);
    Fe canon = hf_from_mont(hf, c);
    for (int d = 0; d < 4; d++) {
        Fe pw = {{1, 0, 0, 0}};   // 2^(64d+96) mod p by doubling
        for (int i = 0; i < 64 * d + 96; i++) {
            uint64_t cy = hf_add_raw(pw.v, pw.v, pw.v);
            if (cy || hf_cmp(pw.v, hf.p) >= 0) hf_sub_raw(pw.v, pw.v, hf.p);
        }
        Fe t = hf_from_mont(hf, hf_mul(hf, hf_to_mont(hf, canon), hf_to_mont(hf, pw)));
        memcpy(tab[d], t.v, 32);
    }
}
extern "C" void emu_const_dot(int field, int n_terms, uint32_t* r, const uint32_t* a, const uint32_t* c_mont, size_t n) {
    const HostField* hf = host_field(field);
    for (size_t i = 0; i < n; i++) {
        uint32_t tab[4][4][8];
        for (int j = 0; j < n_terms; j++) build_digit_table(*hf, c_mont + (i * n_terms + j) * 8, tab[j]);
        const uint32_t(*ap)[8] = reinterpret_cast<const uint32_t(*)[8]>(a + i * n_terms * 8);
#define PSD_CASE(N)                                                                  \
    case N:                                                                          \
        if (field == 0) const_dot<Bn254Fr, N>(r + 8 * i, ap, [&](int j, int d, uint32_t* t) { memcpy(t, tab[j][d], 32); });    \
        else const_dot<Bls12381Fr, N>(r + 8 * i, ap, [&](int j, int d, uint32_t* t) { memcpy(t, tab[j][d], 32); });            \
        break;
        switch (n_terms) { PSD_CASE(1) PSD_CASE(2) PSD_CASE(3) PSD_CASE(4) }
#undef PSD_CASE
    }
}

// ---- fp52.cuh primitives (FP64-pipe products), under round-toward-zero as on the device -------------
extern "C" void emu_f52_prod(const uint64_t* a, const uint64_t* b, uint64_t* hi, uint64_t* lo, size_t n) {
    const int old_round = fegetround();
    fesetround(FE_TOWARDZERO);
    for (size_t i = 0; i < n; i++) {
        uint64_t h, l;
        f52_prod(h, l, (double)a[i], (double)b[i], 20282409603651670423947251286016.0);
        hi[i] = h - kMagicHi;
        lo[i] = l - kMagicLo;
    }
    fesetround(old_round);
}
// 8 x 32-bit words -> five limbs -> 8 x 32-bit words
extern "C" void emu_f52_unpack_pack(const uint32_t* x, uint64_t* limbs, uint32_t* back, size_t n) {
    for (size_t i = 0; i < n; i++) {
        uint64_t m[5];
        f52_unpack(m, x + 8 * i);
        for (int k = 0; k < 5; k++) limbs[5 * i + k] = m[k] - kMagicLo;
        uint64_t plain[5];
        for (int k = 0; k < 5; k++) plain[k] = m[k] & kMask52;
        f52_pack(back + 8 * i, plain);
    }
}


// ---- the numeric bounds behind variant 5, evaluated on the product's own image (optionally damaged) --------
// corrupt: 0 = as built; 1 = one table limb set to 2^52; 2 = every table limb 2^52 - 1 ("all ones": entries
// >= p); 3 = wrong modulus limb; 4 = every lane-update entry replaced by p - 1 (still residues: must pass);
// 5 = one additive constant's limb pushed past 2^52.  Returns 1 if the bounds hold, 0 if not, -1 if the
// variant does not apply to the parameter set; `why` receives the first failed bound.
extern "C" int emu_fp52_bounds(int field, uint32_t rate, uint32_t capacity, uint32_t rf, uint32_t rp, const uint64_t* ark_canon,
                               const uint64_t* mds_canon, int corrupt, char* why, size_t why_len) {
    const HostField* hf = host_field(field);
    const uint32_t t = rate + capacity, rounds = rf + rp;
    std::vector<Fe> ark(rounds * t), mds(t * t);
    for (size_t i = 0; i < ark.size(); i++) ark[i] = hf_to_mont(*hf, reinterpret_cast<const Fe*>(ark_canon)[i]);
    for (size_t i = 0; i < mds.size(); i++) mds[i] = hf_to_mont(*hf, reinterpret_cast<const Fe*>(mds_canon)[i]);
    std::vector<uint32_t> img;
    PermDesc d{};
    append_fp52_tables(*hf, t, rf, rp, ark, mds, d, img);
    if (!d.fp_ok) return -1;
    auto set_dbl = [&img](uint32_t off8, double x) {
        uint64_t q;
        memcpy(&q, &x, 8);
        img[2 * off8] = (uint32_t)q;
        img[2 * off8 + 1] = (uint32_t)(q >> 32);
    };
    const uint32_t n = t - 1;
    if (corrupt == 1) set_dbl(d.fp_mat_off + 7, 4503599627370496.0);
    if (corrupt == 2)
        for (uint32_t o = d.fp_mat_off; o < d.fp_mat_off + t * t * 25; o++) set_dbl(o, 4503599627370495.0);
    if (corrupt == 3) set_dbl(d.fp_consts_off + 2, 12345.0);
    if (corrupt == 4 && rp) {
        uint64_t pm1[4];
        const uint64_t one[4] = {1, 0, 0, 0};
        hf_sub_raw(pm1, hf->p, one);
        for (uint32_t i = 0; i < rp; i++)
            for (uint32_t j = 0; j < n; j++)
                for (int e = 0; e < 5; e++)
                    for (int l = 0; l < 5; l++) {
                        const int bit = 52 * l, w = bit / 64, sh = bit % 64;
                        uint64_t v = pm1[w] >> sh;
                        if (sh > 12 && w + 1 < 4) v |= pm1[w + 1] << (64 - sh);
                        set_dbl(d.fp_part_off + (i * 2 * n + n + j) * 25 + e * 5 + l, (double)(v & ((1ull << 52) - 1)));
                    }
    }
    if (corrupt == 5) {
        const uint32_t o = d.fp_k_off + 2;
        uint64_t q = (uint64_t)img[2 * o] | ((uint64_t)img[2 * o + 1] << 32);
        q += 1ull << 52;
        img[2 * o] = (uint32_t)q;
        img[2 * o + 1] = (uint32_t)(q >> 32);
    }
    std::string reason;
    const bool ok = fp52_image_bounds_ok(*hf, t, rf, rp, d, img, &reason);
    if (why && why_len) snprintf(why, why_len, "%s", reason.c_str());
    return ok ? 1 : 0;
}

// ---- host model of VARIANT 6 (tensor-core linear layer): the product's own tables (tc_tables.cpp) driven per sponge
// by tc_host_row -- the same column sums, sweep and Montgomery row the device computes with mma.sync -- and the
// device's integer S-box code.  Checks the table construction, the constant ordering and the bounds without a GPU.
#include "../../sponge_b200/csrc/tc_tables.hpp"

template <class F, int T, int ALPHA>
static void emu_perm_tc_one(uint32_t (&st)[T][8], const PermDesc& d, const TcDesc& td, const uint8_t* tab, const HostField& hf) {
    const uint32_t kb = kTcKBytes;
    for (int j = 0; j < T; j++) {
        uint32_t c[8];
        cload8(c, d.ark_off + j * 8);
        fp_add<F>(st[j], st[j], c);
    }
    const uint32_t half = d.half_full, rp = d.partial;
    const uint8_t* kp = tab + td.k_off;
    auto full_round = [&](const uint8_t* mat) {
        uint32_t y[T][8];
        for (int j = 0; j < T; j++) {
            tc_sbox<F, ALPHA>(st[j]);
            memcpy(y[j], st[j], 32);
        }
        for (int i = 0; i < T; i++) {
            const uint8_t* frags[T];
            for (int j = 0; j < T; j++) frags[j] = mat + (i * T + j) * kTcFragBytes;
            tc_host_row(hf, &y[0][0], T, frags, kp + i * kb, st[i]);
        }
        kp += T * kb;
    };
    for (uint32_t r = 0; r < half; r++) full_round(tab + ((r + 1 == half && rp) ? td.pre_off : td.mat_off));
    if (rp) {
        const uint8_t* part = tab + td.part_off;
        const uint8_t* pair = tab + td.pair_off;
        const uint32_t round_k = T * kb, round_part = 2 * (T - 1) * kTcFragBytes;
        // one round on its own: row 0, then every lane <- 1 lane + v y
        auto single = [&]() {
            tc_sbox<F, ALPHA>(st[0]);
            uint32_t terms[T][8];
            const uint8_t* frags[T];
            memcpy(terms[0], st[0], 32);
            frags[0] = tab + td.mat_off;
            for (int j = 1; j < T; j++) {
                memcpy(terms[j], st[j], 32);
                frags[j] = part + (j - 1) * kTcFragBytes;
            }
            uint32_t n0[8];
            tc_host_row(hf, &terms[0][0], T, frags, kp, n0);
            for (int j = 1; j < T; j++) {
                uint32_t two[2][8];
                memcpy(two[0], st[0], 32);   // y
                memcpy(two[1], st[j], 32);   // the lane
                const uint8_t* fr[2] = {part + (T - 1 + j - 1) * kTcFragBytes, tab + td.ident_off};
                tc_host_row(hf, &two[0][0], 2, fr, kp + j * kb, st[j]);
            }
            memcpy(st[0], n0, 32);
            kp += round_k;
            part += round_part;
        };
        if (rp & 1) single();
        // pairs (i, i+1): the second row 0 takes the lanes as they were before the pair plus c_q y0; the lanes are
        // brought up to date once per pair
        for (uint32_t q = 0; q < rp / 2; q++) {
            uint32_t y0[8], y1[8], x1[8], x2[8];
            tc_sbox<F, ALPHA>(st[0]);
            memcpy(y0, st[0], 32);
            {
                uint32_t terms[T][8];
                const uint8_t* frags[T];
                memcpy(terms[0], y0, 32);
                frags[0] = tab + td.mat_off;
                for (int j = 1; j < T; j++) {
                    memcpy(terms[j], st[j], 32);
                    frags[j] = part + (j - 1) * kTcFragBytes;
                }
                tc_host_row(hf, &terms[0][0], T, frags, kp, x1);
            }
            memcpy(y1, x1, 32);
            tc_sbox<F, ALPHA>(y1);
            {
                uint32_t terms[T + 1][8];
                const uint8_t* frags[T + 1];
                memcpy(terms[0], y1, 32);
                frags[0] = tab + td.mat_off;
                memcpy(terms[1], y0, 32);
                frags[1] = pair;
                for (int j = 1; j < T; j++) {
                    memcpy(terms[j + 1], st[j], 32);
                    frags[j + 1] = part + round_part + (j - 1) * kTcFragBytes;
                }
                tc_host_row(hf, &terms[0][0], T + 1, frags, kp + round_k, x2);
            }
            for (int j = 1; j < T; j++) {
                uint32_t three[3][8];
                memcpy(three[0], y1, 32);
                memcpy(three[1], y0, 32);
                memcpy(three[2], st[j], 32);
                const uint8_t* fr[3] = {part + round_part + (T - 1 + j - 1) * kTcFragBytes, part + (T - 1 + j - 1) * kTcFragBytes,
                                        tab + td.ident_off};
                tc_host_row(hf, &three[0][0], 3, fr, kp + round_k + j * kb, st[j]);
            }
            memcpy(st[0], x2, 32);
            kp += 2 * round_k;
            part += 2 * round_part;
            pair += kTcFragBytes;
        }
    }
    for (uint32_t r = half + rp; r < d.total; r++) full_round(tab + td.mat_off);
    for (int j = 0; j < T; j++) {
        uint32_t x[8];
        memcpy(x, st[j], 32);
        cond_sub_p<F>(st[j], x);
    }
}

// ---- host model of VARIANT 7: 64-byte unreduced S-box values against the stacked 64-row tables of layout 1, every
// reduction (the squarings' too) as a table contraction + sweep, exactly the sequence of poseidon_tc5.cuh.
template <class F, int ALPHA>
static void emu_tc5_pow(uint32_t* wide16, const uint32_t* x, const uint8_t* red, const HostField& hf) {
    uint32_t y[8], w[16];
    memcpy(y, x, 32);
    for (int q = 0; q < (ALPHA == 5 ? 2 : 4); q++) {
        sqr_wide(w, y);
        const Tc5HostTerm term{w, 64, red, 1, 0};
        tc5_host_row(hf, &term, 1, nullptr, y);
    }
    mul_wide(wide16, y, x);
}

template <class F, int T, int ALPHA>
static void emu_perm_tc5_one(uint32_t (&st)[T][8], const PermDesc& d, const TcDesc& td, const uint8_t* tab, const HostField& hf) {
    constexpr uint32_t N = T - 1, kStack = 2 * T * 1024, kRoundPart = N * 1024 + kStack, kb = kTc5KBytes;
    for (int j = 0; j < T; j++) {
        uint32_t c[8];
        cload8(c, d.ark_off + j * 8);
        fp_add<F>(st[j], st[j], c);
    }
    const uint32_t half = d.half_full, rp = d.partial;
    const uint8_t* kp = tab + td.k_off;
    const uint8_t* red = tab + td.red_off;
    const uint8_t* m00 = tab + td.m00_off;
    const uint8_t* ident = tab + td.ident_off;
    auto full_round = [&](const uint8_t* mat) {
        uint32_t y[T][16];
        for (int j = 0; j < T; j++) emu_tc5_pow<F, ALPHA>(y[j], st[j], red, hf);
        for (int i = 0; i < T; i++) {
            Tc5HostTerm terms[T];
            for (int j = 0; j < T; j++) terms[j] = Tc5HostTerm{y[j], 64, mat + j * kStack, (uint32_t)T, (uint32_t)i};
            tc5_host_row(hf, terms, T, kp + i * kb, st[i], T > 4);
        }
        kp += T * kb;
    };
    for (uint32_t r = 0; r < half; r++) full_round(tab + ((r + 1 == half && rp) ? td.pre_off : td.mat_off));
    if (rp) {
        const uint8_t* part = tab + td.part_off;   // per round: w [N KiB], stack (M00, v) [2 T KiB]
        const uint8_t* pair = tab + td.pair_off;   // per pair: stack (c_q, v(i))
        if (rp & 1) {
            uint32_t y0[16], n0[8];
            emu_tc5_pow<F, ALPHA>(y0, st[0], red, hf);
            const uint8_t* stack = part + N * 1024;
            {
                Tc5HostTerm terms[T];
                terms[0] = Tc5HostTerm{y0, 64, stack, (uint32_t)T, 0};
                for (int j = 1; j < T; j++) terms[j] = Tc5HostTerm{st[j], 32, part + (j - 1) * 1024, 1, 0};
                tc5_host_row(hf, terms, T, kp, n0);
            }
            for (int j = 1; j < T; j++) {
                uint32_t lane[8];
                memcpy(lane, st[j], 32);
                const Tc5HostTerm terms[2] = {Tc5HostTerm{y0, 64, stack, (uint32_t)T, (uint32_t)j}, Tc5HostTerm{lane, 32, ident, 1, 0}};
                tc5_host_row(hf, terms, 2, kp + j * kb, st[j]);
            }
            memcpy(st[0], n0, 32);
            kp += T * kb;
            part += kRoundPart;
        }
        for (uint32_t q = 0; q < rp / 2; q++) {
            const bool last = q + 1 == rp / 2;
            uint32_t y0[16], y1[16], x1[8], x2[8];
            emu_tc5_pow<F, ALPHA>(y0, st[0], red, hf);
            {
                Tc5HostTerm terms[T];
                terms[0] = Tc5HostTerm{y0, 64, m00, 1, 0};
                for (int j = 1; j < T; j++) terms[j] = Tc5HostTerm{st[j], 32, part + (j - 1) * 1024, 1, 0};
                tc5_host_row(hf, terms, T, kp, x1);
            }
            emu_tc5_pow<F, ALPHA>(y1, x1, red, hf);
            const uint8_t* wb = part + kRoundPart;
            const uint8_t* by1 = wb + N * 1024;
            {
                Tc5HostTerm terms[T + 1];
                terms[0] = Tc5HostTerm{y1, 64, by1, (uint32_t)T, 0};
                terms[1] = Tc5HostTerm{y0, 64, pair, (uint32_t)T, 0};
                for (int j = 1; j < T; j++) terms[j + 1] = Tc5HostTerm{st[j], 32, wb + (j - 1) * 1024, 1, 0};
                tc5_host_row(hf, terms, T + 1, kp + T * kb, x2);
            }
            for (int j = 1; j < T; j++) {
                uint32_t lane[8];
                memcpy(lane, st[j], 32);
                const Tc5HostTerm terms[3] = {Tc5HostTerm{y1, 64, by1, (uint32_t)T, (uint32_t)j},
                                              Tc5HostTerm{y0, 64, pair, (uint32_t)T, (uint32_t)j}, Tc5HostTerm{lane, 32, ident, 1, 0}};
                // the device adds the lanes' constant only behind the last partial round; before that it is zero
                tc5_host_row(hf, terms, 3, last ? kp + T * kb + j * kb : nullptr, st[j]);
            }
            memcpy(st[0], x2, 32);
            kp += 2 * T * kb;
            part += 2 * kRoundPart;
            pair += kStack;
        }
    }
    for (uint32_t r = half + rp; r < d.total; r++) full_round(tab + td.mat_off);
    for (int j = 0; j < T; j++) {
        uint32_t x[8];
        memcpy(x, st[j], 32);
        cond_sub_p<F>(st[j], x);
    }
}

template <class F, int T>
static int emu_perm_tc5_a(uint64_t alpha, const PermDesc& d, const TcDesc& td, const uint8_t* tab, const HostField& hf, uint32_t* states,
                          size_t n) {
    for (size_t i = 0; i < n; i++) {
        uint32_t s[T][8];
        memcpy(s, states + i * T * 8, sizeof s);
        if (alpha == 5) emu_perm_tc5_one<F, T, 5>(s, d, td, tab, hf);
        else if (alpha == 17) emu_perm_tc5_one<F, T, 17>(s, d, td, tab, hf);
        else return 1;
        memcpy(states + i * T * 8, s, sizeof s);
    }
    return 0;
}

// the unreduced products on their own: out (16 words) = a * b, or a^2 when b is null
extern "C" void emu_mul_wide(const uint32_t* a, const uint32_t* b, uint32_t* out, size_t n) {
    for (size_t i = 0; i < n; i++) {
        if (b) mul_wide(out + 16 * i, a + 8 * i, b + 8 * i);
        else sqr_wide(out + 16 * i, a + 8 * i);
    }
}

template <class F, int T>
static int emu_perm_tc_a(uint64_t alpha, const PermDesc& d, const TcDesc& td, const uint8_t* tab, const HostField& hf, uint32_t* states,
                         size_t n) {
    for (size_t i = 0; i < n; i++) {
        uint32_t s[T][8];
        memcpy(s, states + i * T * 8, sizeof s);
        if (alpha == 5) emu_perm_tc_one<F, T, 5>(s, d, td, tab, hf);
        else if (alpha == 17) emu_perm_tc_one<F, T, 17>(s, d, td, tab, hf);
        else return 1;
        memcpy(states + i * T * 8, s, sizeof s);
    }
    return 0;
}

// returns 0 ok, 1 if variant 6 does not apply to the parameter set, -1 unsupported width
extern "C" int emu_permute_tc_layout(int field, uint32_t rate, uint32_t capacity, uint32_t rf, uint32_t rp, uint64_t alpha,
                                     const uint64_t* ark_canon, const uint64_t* mds_canon, uint32_t* states, size_t n, uint32_t layout);
extern "C" int emu_permute_tc(int field, uint32_t rate, uint32_t capacity, uint32_t rf, uint32_t rp, uint64_t alpha,
                              const uint64_t* ark_canon, const uint64_t* mds_canon, uint32_t* states, size_t n) {
    return emu_permute_tc_layout(field, rate, capacity, rf, rp, alpha, ark_canon, mds_canon, states, n, 0);
}
extern "C" int emu_permute_tc_layout(int field, uint32_t rate, uint32_t capacity, uint32_t rf, uint32_t rp, uint64_t alpha,
                                     const uint64_t* ark_canon, const uint64_t* mds_canon, uint32_t* states, size_t n, uint32_t layout) {
    const HostField* hf = host_field(field);
    const uint32_t t = rate + capacity, rounds = rf + rp;
    std::vector<Fe> ark(rounds * t), mds(t * t);
    for (size_t i = 0; i < ark.size(); i++) ark[i] = hf_to_mont(*hf, reinterpret_cast<const Fe*>(ark_canon)[i]);
    for (size_t i = 0; i < mds.size(); i++) mds[i] = hf_to_mont(*hf, reinterpret_cast<const Fe*>(mds_canon)[i]);
    TcDesc td{};
    std::vector<uint8_t> tab;
    build_tc_tables(*hf, t, rf, rp, ark, mds, td, tab, layout);
    if (!td.ok) return 1;
    PermDesc d{};
    d.ark_off = 0;
    for (uint32_t j = 0; j < t; j++)
        for (int k = 0; k < 4; k++) {
            c_pool[j * 8 + 2 * k] = (uint32_t)ark[j].v[k];
            c_pool[j * 8 + 2 * k + 1] = (uint32_t)(ark[j].v[k] >> 32);
        }
    d.half_full = rf / 2; d.partial = rp; d.total = rounds; d.rate = rate; d.capacity = capacity;
    if (layout == 1) {
        if (field == 0) {
            switch (t) {
                case 2: return emu_perm_tc5_a<Bn254Fr, 2>(alpha, d, td, tab.data(), *hf, states, n);
                case 3: return emu_perm_tc5_a<Bn254Fr, 3>(alpha, d, td, tab.data(), *hf, states, n);
                case 4: return emu_perm_tc5_a<Bn254Fr, 4>(alpha, d, td, tab.data(), *hf, states, n);
                case 5: return emu_perm_tc5_a<Bn254Fr, 5>(alpha, d, td, tab.data(), *hf, states, n);
            }
        } else {
            switch (t) {
                case 2: return emu_perm_tc5_a<Bls12381Fr, 2>(alpha, d, td, tab.data(), *hf, states, n);
                case 3: return emu_perm_tc5_a<Bls12381Fr, 3>(alpha, d, td, tab.data(), *hf, states, n);
                case 4: return emu_perm_tc5_a<Bls12381Fr, 4>(alpha, d, td, tab.data(), *hf, states, n);
                case 5: return emu_perm_tc5_a<Bls12381Fr, 5>(alpha, d, td, tab.data(), *hf, states, n);
            }
        }
        return -1;
    }
    if (field == 0) {
        switch (t) {
            case 2: return emu_perm_tc_a<Bn254Fr, 2>(alpha, d, td, tab.data(), *hf, states, n);
            case 3: return emu_perm_tc_a<Bn254Fr, 3>(alpha, d, td, tab.data(), *hf, states, n);
            case 4: return emu_perm_tc_a<Bn254Fr, 4>(alpha, d, td, tab.data(), *hf, states, n);
            case 5: return emu_perm_tc_a<Bn254Fr, 5>(alpha, d, td, tab.data(), *hf, states, n);
        }
    } else {
        switch (t) {
            case 2: return emu_perm_tc_a<Bls12381Fr, 2>(alpha, d, td, tab.data(), *hf, states, n);
            case 3: return emu_perm_tc_a<Bls12381Fr, 3>(alpha, d, td, tab.data(), *hf, states, n);
            case 4: return emu_perm_tc_a<Bls12381Fr, 4>(alpha, d, td, tab.data(), *hf, states, n);
            case 5: return emu_perm_tc_a<Bls12381Fr, 5>(alpha, d, td, tab.data(), *hf, states, n);
        }
    }
    return -1;
}
