// Host-only Poseidon parameter generation: field tables, Grain LFSR, Cauchy MDS.
// Replaces /root/reference/src/poseidon/grain_lfsr.rs:15-189 and
// /root/reference/src/poseidon/traits.rs:69-146 (one-time, cold, serial set-up code).
#include "params.hpp"

#include <mutex>
#include <vector>

namespace psd {

// ---------------------------------------------------------------------------------------------
// Field tables
// ---------------------------------------------------------------------------------------------
static HostField g_fields[2];
static std::once_flag g_fields_once;

static void init_field(HostField& f, int id, const uint64_t p[4], int bits) {
    memcpy(f.p, p, 32);
    f.bits = bits;
    f.id = id;
    // -p^-1 mod 2^64 by Newton iteration (p odd)
    uint64_t inv = 1;
    for (int i = 0; i < 6; i++) inv *= 2 - p[0] * inv;
    f.np0 = (uint64_t)0 - inv;
    // R mod p and R^2 mod p by repeated doubling of 1
    Fe x = {{1, 0, 0, 0}};
    for (int i = 0; i < 512; i++) {
        uint64_t c = hf_add_raw(x.v, x.v, x.v);
        if (c || hf_cmp(x.v, f.p) >= 0) hf_sub_raw(x.v, x.v, f.p);
        if (i == 255) memcpy(f.one, x.v, 32);
    }
    memcpy(f.r2, x.v, 32);
}

const HostField* host_field(int field_id) {
    std::call_once(g_fields_once, [] {
        // BN254 Fr (alt_bn128 scalar field) -- not in the reference; BASELINE.json configs 1-3
        const uint64_t bn[4] = {0x43e1f593f0000001ull, 0x2833e84879b97091ull, 0xb85045b68181585dull,
                                0x30644e72e131a029ull};
        // BLS12-381 Fr -- /root/reference/src/test.rs:6
        const uint64_t bls[4] = {0xffffffff00000001ull, 0x53bda402fffe5bfeull, 0x3339d80809a1d805ull,
                                 0x73eda753299d7d48ull};
        init_field(g_fields[0], 0, bn, 254);
        init_field(g_fields[1], 1, bls, 255);
    });
    if (field_id < 0 || field_id > 1) return nullptr;
    return &g_fields[field_id];
}

// ---------------------------------------------------------------------------------------------
// Grain LFSR in self-shrinking mode (grain_lfsr.rs)
// ---------------------------------------------------------------------------------------------
namespace {

class Grain {
   public:
    Grain(uint64_t prime_bits, uint64_t t, uint64_t rf, uint64_t rp) : nbits_(prime_bits), pos_(0) {
        // 80-bit seed, fields written most-significant bit first (grain_lfsr.rs:22-73):
        // [0,1]=field type (01: prime), [2..5]=S-box (0000: x^alpha), [6..17]=n, [18..29]=t,
        // [30..39]=R_F, [40..49]=R_P, [50..79]=1
        for (int i = 0; i < 80; i++) bit_[i] = 0;
        bit_[1] = 1;
        put(6, 12, prime_bits);
        put(18, 12, t);
        put(30, 10, rf);
        put(40, 10, rp);
        for (int i = 50; i < 80; i++) bit_[i] = 1;
        for (int i = 0; i < 160; i++) step();   // grain_lfsr.rs:176-188
    }

    // One self-shrunk output bit: keep drawing pairs until the first of a pair is 1
    // (grain_lfsr.rs:86-106).
    int next_bit() {
        for (;;) {
            int first = step();
            int second = step();
            if (first) return second;
        }
    }

    // nbits_ output bits assembled most-significant first into a 256-bit LE integer
    // (grain_lfsr.rs:119-123 / 141-155).
    Fe draw() {
        Fe v = {{0, 0, 0, 0}};
        for (uint64_t i = 0; i < nbits_; i++) {
            const int b = next_bit();
            const uint64_t pos = nbits_ - 1 - i;
            v.v[pos / 64] |= (uint64_t)b << (pos % 64);
        }
        return v;
    }

   private:
    void put(int first, int width, uint64_t value) {
        for (int i = 0; i < width; i++) bit_[first + width - 1 - i] = (uint8_t)((value >> i) & 1);
    }
    int step() {   // taps 62,51,38,23,13,0 relative to the ring head (grain_lfsr.rs:162-174)
        const int h = pos_;
        const uint8_t nb = bit_[(h + 62) % 80] ^ bit_[(h + 51) % 80] ^ bit_[(h + 38) % 80] ^
                           bit_[(h + 23) % 80] ^ bit_[(h + 13) % 80] ^ bit_[h];
        bit_[h] = nb;
        pos_ = (h + 1) % 80;
        return nb;
    }
    uint8_t bit_[80];
    uint64_t nbits_;
    int pos_;
};

}  // namespace

int find_poseidon_ark_and_mds(const HostField& f, uint64_t prime_bits, uint32_t rate, uint32_t rf,
                              uint32_t rp, uint32_t skip_matrices, Fe* ark, Fe* mds) {
    if ((int)prime_bits != f.bits) return -1;   // assert_eq!, grain_lfsr.rs:112,136
    const uint32_t t = rate + 1;
    Grain g(prime_bits, t, rf, rp);
    // round constants: rejection sampling (grain_lfsr.rs:108-133)
    for (uint32_t i = 0; i < (rf + rp) * t; i++) {
        Fe v;
        do {
            v = g.draw();
        } while (hf_cmp(v.v, f.p) >= 0);
        ark[i] = v;
    }
    // skipped matrices consume 2t mod-p samples each (traits.rs:125-127)
    for (uint32_t s = 0; s < skip_matrices; s++)
        for (uint32_t i = 0; i < 2 * t; i++) (void)g.draw();
    // Cauchy matrix 1/(x_i + y_j) with x, y sampled mod p (traits.rs:134-143, grain_lfsr.rs:135-159)
    std::vector<Fe> xs(t), ys(t);
    for (uint32_t i = 0; i < t; i++) xs[i] = hf_reduce(f, g.draw());
    for (uint32_t i = 0; i < t; i++) ys[i] = hf_reduce(f, g.draw());
    for (uint32_t i = 0; i < t; i++) {
        for (uint32_t j = 0; j < t; j++) {
            Fe s = hf_add(f, xs[i], ys[j]);
            if (hf_is_zero(s)) return -2;   // the reference would panic on inverse().unwrap()
            Fe inv = hf_inv(f, hf_to_mont(f, s));
            mds[i * t + j] = hf_from_mont(f, inv);
        }
    }
    return 0;
}

// (rate, alpha, full_rounds, partial_rounds, skip_matrices) -- test.rs:13-32 for BLS12-381 Fr.
bool default_entry(int field_id, uint32_t rate, bool weights, DefaultEntry* e) {
    if (rate < 2 || rate > 8) return false;
    if (field_id == 1) {
        static const uint32_t rp_c[7] = {31, 56, 56, 57, 57, 57, 57};
        e->rate = rate;
        e->skip_matrices = 0;
        e->full_rounds = 8;
        if (weights) {
            e->alpha = 257;
            e->partial_rounds = 13;
        } else {
            e->alpha = rate == 2 ? 17 : 5;
            e->partial_rounds = rp_c[rate - 2];
        }
        return true;
    }
    if (field_id == 0 && !weights) {
        // BN254 Fr is not in the reference; round numbers for rates 2 and 4 are BASELINE.json's
        // configs (8/57 and 8/60, alpha 5); the other rates use the circomlib table for t=rate+1.
        static const uint32_t rp_bn[7] = {57, 56, 60, 60, 63, 64, 63};
        e->rate = rate;
        e->alpha = 5;
        e->full_rounds = 8;
        e->partial_rounds = rp_bn[rate - 2];
        e->skip_matrices = 0;
        return true;
    }
    return false;
}

}  // namespace psd
