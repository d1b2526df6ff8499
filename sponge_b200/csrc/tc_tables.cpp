// Host-side (one-time) tables for the tensor-core linear layer (poseidon_tc.cuh).
//
// Everything is derived from the regrouped constants of opt_constants.cpp (M_pre, the sparse rows (M00, w), the
// lane coefficients v, the per-round constants k_i and post_ark), exactly as the FP64-pipe variant does
// (fp52_tables.cpp).  A constant c (true value; the state holds Montgomery residues, and (xR) c = (xc) R) becomes the
// 32 x 32 byte matrix Tab_c[k][col] = byte col of (c 256^k 2^32 mod p); the 2^32 is divided out again by the one
// Montgomery row of the sweep.  An additive constant K (a Montgomery residue) becomes the bytes of K 2^32 mod p,
// the start values of the accumulators.
#include "tc_tables.hpp"

#include <string.h>

namespace psd {

namespace {

struct Pow2m {
    const HostField& f;
    Fe two;
    explicit Pow2m(const HostField& hf) : f(hf) {
        Fe t = {{2, 0, 0, 0}};
        two = hf_to_mont(f, t);
    }
    Fe operator()(uint64_t e) const { return hf_pow_u64(f, two, e); }   // Montgomery form of 2^e mod p
};

Fe fe_at(const std::vector<uint32_t>& img, uint32_t word_off) {
    Fe x;
    for (int k = 0; k < 4; k++) x.v[k] = (uint64_t)img[word_off + 2 * k] | ((uint64_t)img[word_off + 2 * k + 1] << 32);
    return x;
}

inline uint8_t byte_of(const Fe& x, int b) { return (uint8_t)(x.v[b / 8] >> (8 * (b % 8))); }

// fragment block of the constant with Montgomery form c_mont: for n-tile nt, lane L = 4 g + tg:
//   b0 = Tab[4 tg + 0..3][8 nt + g],  b1 = Tab[16 + 4 tg + 0..3][8 nt + g]      (lowest byte first)
void push_frag(std::vector<uint8_t>& img, const HostField& f, const Fe pow_mont[32], const Fe& c_mont, uint32_t layout) {
    Fe tab[32];   // tab[k] = c 256^k 2^32 mod p, canonical
    for (int k = 0; k < 32; k++) tab[k] = hf_from_mont(f, hf_mul(f, c_mont, pow_mont[k]));
    if (layout == 1) {
        // B[n = column][k] for tcgen05.mma, K-major, no swizzle: 8-row x 16-byte core matrices, rows contiguous
        for (int kk = 0; kk < 2; kk++)
            for (int col = 0; col < 32; col++)
                for (int i = 0; i < 16; i++) img.push_back(byte_of(tab[16 * kk + i], col));
        return;
    }
    for (int nt = 0; nt < 4; nt++)
        for (int lane = 0; lane < 32; lane++) {
            const int g = lane >> 2, tg = lane & 3, col = 8 * nt + g;
            for (int i = 0; i < 4; i++) img.push_back(byte_of(tab[4 * tg + i], col));
            for (int i = 0; i < 4; i++) img.push_back(byte_of(tab[16 + 4 * tg + i], col));
        }
}

// accumulator start values of the additive constant with Montgomery residue k_mont: bytes of K' = K 2^32 mod p;
// for n-tile nt and column pair tg four little-endian s32: K'[8 nt + 2 tg], K'[8 nt + 2 tg + 1], and the same
// pair again (the C operand of mma.m16n8k32 holds two rows of the pair)
void push_kfrag(std::vector<uint8_t>& img, const HostField& f, const Fe& pow32_mont, const Fe& k_mont, uint32_t layout) {
    // k_mont IS the integer to be added to the state (a residue below p); K' = k_mont * 2^32 mod p as an integer:
    // hf_mul(a, b) = a b / R, so with b = 2^32 R mod p (the Montgomery form of 2^32) the product is a 2^32 mod p.
    const Fe kp = hf_mul(f, k_mont, pow32_mont);
    if (layout == 1) {
        for (int b = 0; b < 32; b++) img.push_back(byte_of(kp, b));
        return;
    }
    for (int nt = 0; nt < 4; nt++)
        for (int tg = 0; tg < 4; tg++)
            for (int e = 0; e < 4; e++) {
                const uint32_t v = byte_of(kp, 8 * nt + 2 * tg + (e & 1));
                for (int b = 0; b < 4; b++) img.push_back((uint8_t)(v >> (8 * b)));
            }
}

// Layout 1, operands that are UNREDUCED products of two Montgomery residues (64 bytes, value = x y R^2 for the true x y):
// the table of the constant c has 64 rows, row k = bytes of (c R^-1 256^k 2^32 mod p), so that the sweep's Montgomery row
// leaves c x y R.  Several constants that multiply the SAME operand are stacked along N -- one tcgen05.mma then fills
// 32 columns per constant: [16-byte K chunk 0..3][constant 0..n-1][column 0..31][16 bytes], 2 n KiB.
void push_stack_wide(std::vector<uint8_t>& img, const HostField& f, const Fe pow_mont[64], const std::vector<Fe>& c_mont) {
    std::vector<std::vector<Fe>> tab(c_mont.size(), std::vector<Fe>(64));
    for (size_t b = 0; b < c_mont.size(); b++)
        for (int k = 0; k < 64; k++) tab[b][k] = hf_from_mont(f, hf_from_mont(f, hf_mul(f, c_mont[b], pow_mont[k])));
    for (int chunk = 0; chunk < 4; chunk++)
        for (size_t b = 0; b < c_mont.size(); b++)
            for (int col = 0; col < 32; col++)
                for (int i = 0; i < 16; i++) img.push_back(byte_of(tab[b][16 * chunk + i], col));
}

}  // namespace

bool tc_bounds_ok(const HostField& f, uint32_t t, std::string* why, uint32_t layout) {
    auto no = [&](const char* m) {
        if (why) *why = m;
        return false;
    };
    if (t < 2 || t > 5) return no("width outside [2,5]");
    // column sums, in units of 32 products of two bytes.  Layout 0: at most t + 1 values of 32 bytes (the second row 0 of
    // a pair of partial rounds) plus one constant byte.  Layout 1: an S-box output is 64 bytes (two units); a full round
    // sums t of them, the second row 0 two of them and t - 1 lanes; the constant is added after the pair-combine.
    // Pair-combining needs sum * 257 < 2^32.
    // At t = 5 the full rounds' sums (ten units) are combined in 64 bits; they only have to fit the s32 accumulators.
    const uint64_t units = layout == 1 ? (2 * t > t + 3 ? 2 * t : t + 3) : t + 1;
    const uint64_t colmax = 32ull * units * 255 * 255 + (layout == 1 ? 0 : 255);
    const uint64_t pair_units = (layout == 1 && t > 4) ? t + 3 : units;
    if ((32ull * pair_units * 255 * 255 + (layout == 1 ? 0 : 255)) * 257 >= (1ull << 32)) return no("column sums too large for the pair-combine");
    if (colmax >= (1ull << 31)) return no("column sums do not fit the accumulators");
    // sweep input V < colmax * 256^32 * 256/255 (+ K' < p in layout 1); result (V + m p) / 2^32 < V / 2^32 + p must stay
    // below 2^256 and, for the lazy S-box (BN254), inside its operand range 2 a + p < 2^256 with the output below 2^256
    long double p = 0;
    for (int k = 3; k >= 0; k--) p = p * 0x1p64L + (long double)f.p[k];
    const long double vmax = (long double)colmax * 0x1p256L * (256.0L / 255.0L) + (layout == 1 ? p : 0.0L);
    const long double res = vmax / 0x1p32L + p;
    if (res >= 0x1p256L) return no("sweep results do not fit 256 bits");
    if (layout == 1) {
        // sqr_wide doubles its operand into eight limbs; the closing conditional subtraction needs a value below 2p
        if (res >= 0x1p255L) return no("sweep results do not fit the squaring's operand range");
        if (res >= 2 * p) return no("sweep results are not below 2p");
        return true;
    }
    if (f.bits == 254) {
        const long double q = p / 0x1p256L;
        long double y = res / p;
        const long double x0 = y;
        for (int sq = 0; sq < 4; sq++) {
            if ((2 * y + 1) * p >= 0x1p256L) return no("a squaring operand leaves the 9-limb window");
            y = 1 + y * y * q;
        }
        if ((y + 1) * p >= 0x1p256L) return no("the closing product's operand is too large");
        if ((1 + y * x0 * q) * p >= 0x1p256L) return no("the lazy S-box output does not fit 256 bits");
    } else if (res >= 2 * p) {
        return no("sweep results are not below 2p");
    }
    return true;
}

void build_tc_tables(const HostField& f, uint32_t t, uint32_t full_rounds, uint32_t partial_rounds,
                     const std::vector<Fe>& ark, const std::vector<Fe>& mds, TcDesc& td, std::vector<uint8_t>& img, uint32_t layout) {
    memset(&td, 0, sizeof td);
    td.layout = layout;
    img.clear();
    const uint32_t half = full_rounds / 2, rp = partial_rounds, n = t - 1, total = full_rounds + rp;
    if (half == 0 || full_rounds - half == 0) return;
    if (!tc_bounds_ok(f, t, nullptr, layout)) return;

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

    const Pow2m pow2(f);
    Fe pow_mont[32];
    for (int k = 0; k < 32; k++) pow_mont[k] = pow2(8 * k + 32);
    const Fe pow32 = pow2(32);
    auto here = [&img]() { return (uint32_t)img.size(); };

    if (layout == 1) {
        // 64-row tables for the S-box outputs, stacked by the operand they multiply (see push_stack_wide)
        Fe pow64[64];
        for (int k = 0; k < 64; k++) pow64[k] = pow2(8 * k + 32);
        auto column = [&](const std::vector<Fe>& m, uint32_t j) {
            std::vector<Fe> c;
            for (uint32_t i = 0; i < t; i++) c.push_back(m[i * t + j]);
            return c;
        };
        td.mat_off = here();   // [T] stacks: column j of the dense matrix (what multiplies S-box output j, for every row)
        for (uint32_t j = 0; j < t; j++) push_stack_wide(img, f, pow64, column(mds, j));
        if (rp) {
            td.pre_off = here();
            for (uint32_t j = 0; j < t; j++) push_stack_wide(img, f, pow64, column(m_pre, j));
        }
        td.ident_off = here();
        push_frag(img, f, pow_mont, hf_one(f), 1);
        td.red_off = here();
        push_stack_wide(img, f, pow64, {hf_one(f)});
        td.m00_off = here();
        push_stack_wide(img, f, pow64, {mds[0]});
        // per partial round: w_1.. (32-row tables, they multiply lanes), then the stack (M00, v_1, ..) that multiplies the
        // round's S-box output: row 0 and every lane in one operand
        td.part_off = here();
        for (uint32_t i = 0; i < rp; i++) {
            for (uint32_t j = 0; j < n; j++) push_frag(img, f, pow_mont, w[i][j], 1);
            std::vector<Fe> c{mds[0]};
            for (uint32_t j = 0; j < n; j++) c.push_back(v[i][j]);
            push_stack_wide(img, f, pow64, c);
        }
        // per pair (i, i+1): the stack (c_q, v_1(i), ..) that multiplies the FIRST round's S-box output in the pair's
        // second step, c_q = sum_j w_j(i+1) v_j(i)
        td.pair_off = here();
        for (uint32_t i = rp & 1; i + 1 < rp; i += 2) {
            Fe c = hf_zero();
            for (uint32_t j = 0; j < n; j++) c = hf_add(f, c, hf_mul(f, w[i + 1][j], v[i][j]));
            std::vector<Fe> cs{c};
            for (uint32_t j = 0; j < n; j++) cs.push_back(v[i][j]);
            push_stack_wide(img, f, pow64, cs);
        }
    } else {
        td.mat_off = here();
        for (uint32_t i = 0; i < t * t; i++) push_frag(img, f, pow_mont, mds[i], layout);
        if (rp) {
            td.pre_off = here();
            for (uint32_t i = 0; i < t * t; i++) push_frag(img, f, pow_mont, m_pre[i], layout);
        }
        td.ident_off = here();
        push_frag(img, f, pow_mont, hf_one(f), layout);
        td.part_off = here();
        for (uint32_t i = 0; i < rp; i++) {
            for (uint32_t j = 0; j < n; j++) push_frag(img, f, pow_mont, w[i][j], layout);
            for (uint32_t j = 0; j < n; j++) push_frag(img, f, pow_mont, v[i][j], layout);
        }
        // pairs of partial rounds (an odd count leaves the FIRST round on its own): c_q = sum_j w_j(i+1) v_j(i)
        td.pair_off = here();
        for (uint32_t i = rp & 1; i + 1 < rp; i += 2) {
            Fe c = hf_zero();
            for (uint32_t j = 0; j < n; j++) c = hf_add(f, c, hf_mul(f, w[i + 1][j], v[i][j]));
            push_frag(img, f, pow_mont, c, layout);
        }
    }
    // additive constants in execution order, t per round: what round r's outputs receive is what round r + 1
    // adds before its S-boxes
    td.k_off = here();
    const Fe zero = hf_zero();
    for (uint32_t r = 0; r < total; r++) {
        const bool full = r < half || r >= half + rp;
        const bool last_partial = !full && r + 1 == half + rp;
        for (uint32_t j = 0; j < t; j++) {
            Fe k = zero;
            if (last_partial) k = post_ark[j];
            else if (!full) k = (j == 0) ? k_lane0[r - half + 1] : zero;
            else if (r + 1 == half && rp) k = (j == 0) ? k_lane0[0] : zero;
            else if (r + 1 < total) k = ark[(size_t)(r + 1) * t + j];
            push_kfrag(img, f, pow32, k, layout);
        }
    }
    while (img.size() % 16) img.push_back(0);
    td.ok = 1;
}

namespace {
// V (nine 32-bit limbs, V[0..8]) -> (V + m p) / 2^32, eight limbs
void host_mont_row(const HostField& f, const uint32_t* v, uint32_t* out8) {
    const uint32_t p32[8] = {(uint32_t)f.p[0], (uint32_t)(f.p[0] >> 32), (uint32_t)f.p[1], (uint32_t)(f.p[1] >> 32),
                             (uint32_t)f.p[2], (uint32_t)(f.p[2] >> 32), (uint32_t)f.p[3], (uint32_t)(f.p[3] >> 32)};
    const uint32_t m = v[0] * (uint32_t)f.np0;
    uint64_t c = 0;
    uint32_t r[10];
    for (int l = 0; l < 9; l++) {
        const unsigned __int128 ss = (unsigned __int128)v[l] + (l < 8 ? (unsigned __int128)m * p32[l] : 0) + c;
        r[l] = (uint32_t)ss;
        c = (uint64_t)(ss >> 32);
    }
    for (int l = 0; l < 8; l++) out8[l] = r[l + 1];
}
}  // namespace

void tc5_host_row(const HostField& f, const Tc5HostTerm* terms, uint32_t n_terms, const uint8_t* kbytes, uint32_t* out8, bool big) {
    uint64_t col[32] = {0};
    for (uint32_t j = 0; j < n_terms; j++) {
        const Tc5HostTerm& t = terms[j];
        for (uint32_t k = 0; k < t.nbytes; k++) {
            const uint32_t byte = (t.value[k / 4] >> (8 * (k % 4))) & 0xff;
            for (int c = 0; c < 32; c++)
                col[c] += (uint64_t)byte * t.table[(k / 16) * (t.stack * 512) + t.block * 512 + c * 16 + (k % 16)];
        }
    }
    // the device pair-combines columns in 32 bits: the sums must leave room for that
    for (int c = 0; c < 32; c++)
        if (big ? col[c] >= (1ull << 31) : col[c] * 257 >= (1ull << 32)) {
            for (int l = 0; l < 8; l++) out8[l] = 0xdeadbeefu;
            return;
        }
    uint32_t v[10] = {0};
    unsigned __int128 carry = 0;
    for (int l = 0; l < 9; l++) {
        unsigned __int128 s = carry;
        for (int b = 0; b < 4; b++)
            if (4 * l + b < 32) s += (unsigned __int128)col[4 * l + b] << (8 * b);
        if (kbytes && l < 8)
            s += (uint32_t)kbytes[4 * l] | ((uint32_t)kbytes[4 * l + 1] << 8) | ((uint32_t)kbytes[4 * l + 2] << 16) |
                 ((uint32_t)kbytes[4 * l + 3] << 24);
        v[l] = (uint32_t)s;
        carry = s >> 32;
    }
    host_mont_row(f, v, out8);
}

void tc_host_row(const HostField& f, const uint32_t* x, uint32_t n_terms, const uint8_t* const* frags, const uint8_t* kfrag,
                 uint32_t* out8) {
    // column sums, read back from the table layout
    uint64_t col[32];
    for (int c = 0; c < 32; c++) {
        const int nt = c / 8, tg = (c % 8) / 2, e = c % 2;
        const uint8_t* q = kfrag + ((nt * 4 + tg) * 4 + e) * 4;
        col[c] = (uint64_t)q[0] | ((uint64_t)q[1] << 8) | ((uint64_t)q[2] << 16) | ((uint64_t)q[3] << 24);
    }
    for (uint32_t j = 0; j < n_terms; j++)
        for (int k = 0; k < 32; k++) {
            const uint32_t byte = (x[j * 8 + k / 4] >> (8 * (k % 4))) & 0xff;
            for (int c = 0; c < 32; c++) {
                const int nt = c / 8, g = c % 8;
                const int tg = (k % 16) / 4, i = k % 4, hi = k / 16;
                const uint8_t tb = frags[j][((nt * 32 + (4 * g + tg)) * 2 + hi) * 4 + i];
                col[c] += (uint64_t)byte * tb;
            }
        }
    // V = sum col[c] 256^c  (nine 32-bit limbs)
    uint32_t v[10] = {0};
    {
        unsigned __int128 carry = 0;
        for (int l = 0; l < 9; l++) {
            unsigned __int128 s = carry;
            for (int b = 0; b < 4; b++)
                if (4 * l + b < 32) s += (unsigned __int128)col[4 * l + b] << (8 * b);
            v[l] = (uint32_t)s;
            carry = s >> 32;
        }
    }
    // one 32-bit Montgomery row
    const uint32_t p32[8] = {(uint32_t)f.p[0], (uint32_t)(f.p[0] >> 32), (uint32_t)f.p[1], (uint32_t)(f.p[1] >> 32),
                             (uint32_t)f.p[2], (uint32_t)(f.p[2] >> 32), (uint32_t)f.p[3], (uint32_t)(f.p[3] >> 32)};
    const uint32_t m = v[0] * (uint32_t)f.np0;
    uint64_t c = 0;
    uint32_t r[10];
    for (int l = 0; l < 9; l++) {
        const unsigned __int128 ss = (unsigned __int128)v[l] + (l < 8 ? (unsigned __int128)m * p32[l] : 0) + c;
        r[l] = (uint32_t)ss;
        c = (uint64_t)(ss >> 32);
    }
    for (int l = 0; l < 8; l++) out8[l] = r[l + 1];
}

}  // namespace psd
