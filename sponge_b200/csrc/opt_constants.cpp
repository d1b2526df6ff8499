// Host-side (one-time) derivation of the constants used by permute_opt (poseidon_core.cuh).
//
// The reference applies the dense MDS matrix M and all t round constants in every round
// (/root/reference/src/poseidon/mod.rs:76-118).  In the partial rounds only lane 0 passes through the
// S-box, so the linear parts of consecutive partial rounds can be regrouped without changing the
// function (the construction described for "optimised Poseidon" in the Poseidon paper's
// supplementary material, restated here for the reference's round order ark -> S-box -> MDS):
//
//  Matrices.  Write A = [[A00, A01],[A10, Â]] with Â the lower-right (t-1)x(t-1) block.  Then
//    A = S · D,   D = diag(1, Â),   S = [[A00, A01·Â^-1],[A10, I]]   (S is "sparse": 2t-1 entries).
//  D fixes lane 0, hence commutes with the partial S-box and can be pushed into the previous
//  round's matrix.  Going backwards from the last partial round:
//    A_RP = M;  A_i = S_i · D_i;  A_{i-1} = D_i · M;  ...;  M_pre = D_1 · M
//  M_pre replaces M in the last full round before the partial rounds.
//
//  Constants.  In the transformed domain partial round i adds c'_i = D_i · c_i.  The part of a
//  constant on lanes 1..t-1 is not touched by the partial S-box, so it is carried forward through
//  S_i and merged into the next round's constant; each partial round is left with ONE constant k_i
//  on lane 0, and what remains after the last partial round is added to the round constants of the
//  first full round that follows (post_ark).
//
// All arithmetic is on Montgomery residues with the host field of host_field.hpp.
#include "opt_constants.hpp"

namespace psd {

namespace {

typedef std::vector<Fe> Mat;   // row-major n x n

// Gauss-Jordan inverse of an n x n matrix; false if singular.
bool mat_inverse(const HostField& f, const Mat& a, int n, Mat& inv) {
    Mat m = a;
    inv.assign((size_t)n * n, hf_zero());
    for (int i = 0; i < n; i++) inv[(size_t)i * n + i] = hf_one(f);
    for (int c = 0; c < n; c++) {
        int piv = -1;
        for (int r = c; r < n; r++) {
            if (!hf_is_zero(m[(size_t)r * n + c])) {
                piv = r;
                break;
            }
        }
        if (piv < 0) return false;
        if (piv != c) {
            for (int k = 0; k < n; k++) {
                std::swap(m[(size_t)piv * n + k], m[(size_t)c * n + k]);
                std::swap(inv[(size_t)piv * n + k], inv[(size_t)c * n + k]);
            }
        }
        const Fe s = hf_inv(f, m[(size_t)c * n + c]);
        for (int k = 0; k < n; k++) {
            m[(size_t)c * n + k] = hf_mul(f, m[(size_t)c * n + k], s);
            inv[(size_t)c * n + k] = hf_mul(f, inv[(size_t)c * n + k], s);
        }
        for (int r = 0; r < n; r++) {
            if (r == c || hf_is_zero(m[(size_t)r * n + c])) continue;
            const Fe fct = m[(size_t)r * n + c];
            for (int k = 0; k < n; k++) {
                m[(size_t)r * n + k] = hf_sub(f, m[(size_t)r * n + k], hf_mul(f, fct, m[(size_t)c * n + k]));
                inv[(size_t)r * n + k] = hf_sub(f, inv[(size_t)r * n + k], hf_mul(f, fct, inv[(size_t)c * n + k]));
            }
        }
    }
    return true;
}

void push_fe(std::vector<uint32_t>& img, const Fe& x) {
    for (int k = 0; k < 4; k++) {
        img.push_back((uint32_t)x.v[k]);
        img.push_back((uint32_t)(x.v[k] >> 32));
    }
}

}  // namespace

void append_optimised_constants(const HostField& f, uint32_t t, uint32_t full_rounds, uint32_t partial_rounds,
                                const std::vector<Fe>& ark, const std::vector<Fe>& mds, PermDesc& d,
                                std::vector<uint32_t>& img) {
    const uint32_t half = full_rounds / 2, rp = partial_rounds, n = t - 1;
    d.opt_ok = 0;
    d.pre_mds_off = d.sparse_off = d.post_ark_off = 0;
    if (rp == 0) {   // nothing to regroup; permute_opt runs full rounds only
        d.opt_ok = 1;
        return;
    }
    // needs a full round before (to absorb M_pre) and after (to absorb the leftover constant)
    if (half == 0 || full_rounds - half == 0) return;

    std::vector<Mat> hats(rp);             // Â_i
    std::vector<std::vector<Fe>> w(rp), v(rp);
    Mat a = mds;                           // A_RP = M
    for (uint32_t ii = rp; ii-- > 0;) {
        Mat hat((size_t)n * n), inv;
        for (uint32_t r = 0; r < n; r++)
            for (uint32_t c = 0; c < n; c++) hat[(size_t)r * n + c] = a[(size_t)(r + 1) * t + (c + 1)];
        if (!mat_inverse(f, hat, (int)n, inv)) return;   // not factorable: keep the dense kernel
        w[ii].resize(n);
        v[ii].resize(n);
        for (uint32_t c = 0; c < n; c++) {
            Fe acc = hf_zero();
            for (uint32_t k = 0; k < n; k++) acc = hf_add(f, acc, hf_mul(f, a[1 + k], inv[(size_t)k * n + c]));
            w[ii][c] = acc;
            v[ii][c] = a[(size_t)(c + 1) * t];
        }
        hats[ii] = hat;
        // A <- diag(1, Â) · M
        Mat next((size_t)t * t);
        for (uint32_t c = 0; c < t; c++) next[c] = mds[c];
        for (uint32_t r = 0; r < n; r++)
            for (uint32_t c = 0; c < t; c++) {
                Fe acc = hf_zero();
                for (uint32_t k = 0; k < n; k++)
                    acc = hf_add(f, acc, hf_mul(f, hat[(size_t)r * n + k], mds[(size_t)(k + 1) * t + c]));
                next[(size_t)(r + 1) * t + c] = acc;
            }
        a = next;
    }
    const Mat& m_pre = a;

    // constants
    std::vector<Fe> g(t, hf_zero()), k_lane0(rp);
    for (uint32_t i = 0; i < rp; i++) {
        const Fe* c = &ark[(size_t)(half + i) * t];
        std::vector<Fe> h(t);
        h[0] = hf_add(f, g[0], c[0]);
        for (uint32_t r = 0; r < n; r++) {
            Fe acc = hf_zero();
            for (uint32_t k = 0; k < n; k++) acc = hf_add(f, acc, hf_mul(f, hats[i][(size_t)r * n + k], c[1 + k]));
            h[1 + r] = hf_add(f, g[1 + r], acc);
        }
        k_lane0[i] = h[0];
        Fe g0 = hf_zero();
        for (uint32_t j = 0; j < n; j++) g0 = hf_add(f, g0, hf_mul(f, w[i][j], h[1 + j]));
        g[0] = g0;
        for (uint32_t j = 0; j < n; j++) g[1 + j] = h[1 + j];
    }

    d.pre_mds_off = (uint32_t)img.size();
    for (const Fe& x : m_pre) push_fe(img, x);
    d.sparse_off = (uint32_t)img.size();
    for (uint32_t i = 0; i < rp; i++) {
        push_fe(img, k_lane0[i]);
        push_fe(img, mds[0]);   // M00, so that the row-0 coefficients (M00, w_1, ..) are contiguous
        for (uint32_t j = 0; j < n; j++) push_fe(img, w[i][j]);
        for (uint32_t j = 0; j < n; j++) push_fe(img, v[i][j]);
    }
    d.post_ark_off = (uint32_t)img.size();
    for (uint32_t j = 0; j < t; j++) push_fe(img, hf_add(f, ark[(size_t)(half + rp) * t + j], g[j]));
    d.opt_ok = 1;
}

namespace {

// tab[dgt] = c * 2^(64 dgt + 96) mod p, canonical, for a Montgomery-form constant c
void push_digit_table(const HostField& f, const Fe& c_mont, const Fe pow_mont[4], std::vector<uint32_t>& img) {
    for (int dgt = 0; dgt < 4; dgt++) {
        // (cR)(XR)/R = cXR -> from_mont -> cX
        push_fe(img, hf_from_mont(f, hf_mul(f, c_mont, pow_mont[dgt])));
    }
}

Fe fe_from_image(const std::vector<uint32_t>& img, uint32_t word_off) {
    Fe x;
    for (int k = 0; k < 4; k++) x.v[k] = (uint64_t)img[word_off + 2 * k] | ((uint64_t)img[word_off + 2 * k + 1] << 32);
    return x;
}

}  // namespace

void append_digit_tables(const HostField& f, uint32_t t, uint32_t full_rounds, uint32_t partial_rounds,
                         const std::vector<Fe>& mds, PermDesc& d, std::vector<uint32_t>& img, bool lanes_only) {
    (void)full_rounds;
    d.ctab_ok = 0;
    d.ct_m00_off = d.ct_mds_off = d.ct_pre_off = d.ct_sparse_off = 0;
    if (!d.opt_ok) return;
    // Montgomery forms of 2^(64 dgt + 96) mod p, by doubling
    Fe pow_mont[4];
    {
        Fe x = {{1, 0, 0, 0}};
        int have = 0;
        for (int dgt = 0; dgt < 4; dgt++) {
            const int want = 64 * dgt + 96;
            for (; have < want; have++) {
                uint64_t c = hf_add_raw(x.v, x.v, x.v);
                if (c || hf_cmp(x.v, f.p) >= 0) hf_sub_raw(x.v, x.v, f.p);
            }
            pow_mont[dgt] = hf_to_mont(f, x);
        }
    }
    const uint32_t rp = partial_rounds, n = t - 1;
    if (lanes_only) {
        if (!rp) return;
        const uint32_t sparse = d.sparse_off;
        std::vector<Fe> vs((size_t)rp * n);
        for (uint32_t r = 0; r < rp; r++)
            for (uint32_t j = 0; j < n; j++) vs[(size_t)r * n + j] = fe_from_image(img, sparse + (r * 2 * t + 2 + n + j) * 8);
        d.ct_sparse_off = (uint32_t)img.size();
        for (const Fe& c : vs) push_digit_table(f, c, pow_mont, img);
        d.ctab_ok = 2;
        return;
    }
    d.ct_m00_off = (uint32_t)img.size();
    push_digit_table(f, mds[0], pow_mont, img);
    d.ct_mds_off = (uint32_t)img.size();
    for (uint32_t i = 0; i < t * t; i++) push_digit_table(f, mds[i], pow_mont, img);
    if (rp) {
        // M_pre and the sparse rows are re-read from the image written by append_optimised_constants
        const uint32_t pre = d.pre_mds_off, sparse = d.sparse_off;
        std::vector<Fe> pre_m(t * t);
        for (uint32_t i = 0; i < t * t; i++) pre_m[i] = fe_from_image(img, pre + i * 8);
        std::vector<Fe> wv((size_t)rp * 2 * n);
        for (uint32_t r = 0; r < rp; r++)
            for (uint32_t j = 0; j < 2 * n; j++) wv[(size_t)r * 2 * n + j] = fe_from_image(img, sparse + (r * 2 * t + 2 + j) * 8);
        d.ct_pre_off = (uint32_t)img.size();
        for (uint32_t i = 0; i < t * t; i++) push_digit_table(f, pre_m[i], pow_mont, img);
        d.ct_sparse_off = (uint32_t)img.size();
        for (const Fe& c : wv) push_digit_table(f, c, pow_mont, img);
    }
    d.ctab_ok = 1;
}

void offset_desc(PermDesc& d, uint32_t pool_off) {
    d.ct_m00_off += pool_off;
    d.ct_mds_off += pool_off;
    d.ct_pre_off += pool_off;
    d.ct_sparse_off += pool_off;
    d.ark_off += pool_off;
    d.mds_off += pool_off;
    d.one_off += pool_off;
    d.r2_off += pool_off;
    d.pre_mds_off += pool_off;
    d.sparse_off += pool_off;
    d.post_ark_off += pool_off;
    // 8-byte units; slots are multiples of 8 words, so pool_off is even
    d.fp_consts_off += pool_off / 2;
    d.fp_mat_off += pool_off / 2;
    d.fp_pre_off += pool_off / 2;
    d.fp_last_off += pool_off / 2;
    d.fp_k_off += pool_off / 2;
    d.fp_part_off += pool_off / 2;
}

}  // namespace psd
