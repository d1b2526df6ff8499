// Host-side 256-bit prime-field arithmetic (4 x 64-bit limbs) used ONLY for one-time set-up:
// parameter generation (Grain LFSR, Cauchy MDS inverse), conversion of constants to Montgomery
// form, and derivation of the optimised round constants.  It never touches sponge states -- the
// permutation itself has no CPU path in this library.
#pragma once
#include <stdint.h>
#include <string.h>

namespace psd {

typedef unsigned __int128 u128;

struct HostField {
    uint64_t p[4];
    uint64_t np0;      // -p^-1 mod 2^64
    uint64_t one[4];   // R mod p
    uint64_t r2[4];    // R^2 mod p
    int bits;
    int id;
};

struct Fe {  // a field element, representation determined by context
    uint64_t v[4];
    bool operator==(const Fe& o) const { return memcmp(v, o.v, 32) == 0; }
};

static inline int hf_cmp(const uint64_t* a, const uint64_t* b) {
    for (int i = 3; i >= 0; i--) {
        if (a[i] != b[i]) return a[i] < b[i] ? -1 : 1;
    }
    return 0;
}
static inline uint64_t hf_add_raw(uint64_t* r, const uint64_t* a, const uint64_t* b) {
    u128 c = 0;
    for (int i = 0; i < 4; i++) {
        c += (u128)a[i] + b[i];
        r[i] = (uint64_t)c;
        c >>= 64;
    }
    return (uint64_t)c;
}
static inline uint64_t hf_sub_raw(uint64_t* r, const uint64_t* a, const uint64_t* b) {
    uint64_t borrow = 0;
    for (int i = 0; i < 4; i++) {
        u128 d = (u128)a[i] - b[i] - borrow;
        r[i] = (uint64_t)d;
        borrow = (uint64_t)(d >> 64) & 1;
    }
    return borrow;
}
static inline bool hf_is_zero(const Fe& a) { return (a.v[0] | a.v[1] | a.v[2] | a.v[3]) == 0; }

static inline Fe hf_add(const HostField& f, const Fe& a, const Fe& b) {
    Fe r;
    uint64_t c = hf_add_raw(r.v, a.v, b.v);
    if (c || hf_cmp(r.v, f.p) >= 0) hf_sub_raw(r.v, r.v, f.p);
    return r;
}
static inline Fe hf_sub(const HostField& f, const Fe& a, const Fe& b) {
    Fe r;
    if (hf_sub_raw(r.v, a.v, b.v)) hf_add_raw(r.v, r.v, f.p);
    return r;
}
static inline Fe hf_neg(const HostField& f, const Fe& a) {
    Fe z = {{0, 0, 0, 0}};
    return hf_sub(f, z, a);
}
// Montgomery product a*b*2^-256 mod p (inputs < p)
static inline Fe hf_mul(const HostField& f, const Fe& a, const Fe& b) {
    uint64_t t[6] = {0, 0, 0, 0, 0, 0};
    for (int i = 0; i < 4; i++) {
        u128 c = 0;
        for (int j = 0; j < 4; j++) {
            c += (u128)a.v[j] * b.v[i] + t[j];
            t[j] = (uint64_t)c;
            c >>= 64;
        }
        c += t[4];
        t[4] = (uint64_t)c;
        t[5] = (uint64_t)(c >> 64);
        const uint64_t m = t[0] * f.np0;
        c = (u128)m * f.p[0] + t[0];
        c >>= 64;
        for (int j = 1; j < 4; j++) {
            c += (u128)m * f.p[j] + t[j];
            t[j - 1] = (uint64_t)c;
            c >>= 64;
        }
        c += t[4];
        t[3] = (uint64_t)c;
        t[4] = t[5] + (uint64_t)(c >> 64);
    }
    Fe r = {{t[0], t[1], t[2], t[3]}};
    if (t[4] || hf_cmp(r.v, f.p) >= 0) hf_sub_raw(r.v, r.v, f.p);
    return r;
}
static inline Fe hf_one(const HostField& f) {
    Fe r;
    memcpy(r.v, f.one, 32);
    return r;
}
static inline Fe hf_zero() { return Fe{{0, 0, 0, 0}}; }
static inline Fe hf_to_mont(const HostField& f, const Fe& a) {
    Fe r2;
    memcpy(r2.v, f.r2, 32);
    return hf_mul(f, a, r2);
}
static inline Fe hf_from_mont(const HostField& f, const Fe& a) {
    Fe one = {{1, 0, 0, 0}};
    return hf_mul(f, a, one);
}
// a^e for a Montgomery residue a and a 256-bit exponent
static inline Fe hf_pow(const HostField& f, const Fe& a, const uint64_t e[4]) {
    Fe acc = hf_one(f);
    for (int i = 255; i >= 0; i--) {
        acc = hf_mul(f, acc, acc);
        if ((e[i / 64] >> (i % 64)) & 1) acc = hf_mul(f, acc, a);
    }
    return acc;
}
static inline Fe hf_pow_u64(const HostField& f, const Fe& a, uint64_t e) {
    uint64_t ee[4] = {e, 0, 0, 0};
    return hf_pow(f, a, ee);
}
// inverse of a non-zero Montgomery residue (Fermat)
static inline Fe hf_inv(const HostField& f, const Fe& a) {
    uint64_t e[4];
    const uint64_t two[4] = {2, 0, 0, 0};
    hf_sub_raw(e, f.p, two);
    return hf_pow(f, a, e);
}
// reduce an arbitrary 256-bit integer mod p (canonical result)
static inline Fe hf_reduce(const HostField& f, const Fe& a) {
    Fe r = a;
    while (hf_cmp(r.v, f.p) >= 0) hf_sub_raw(r.v, r.v, f.p);
    return r;
}

const HostField* host_field(int field_id);   // params.cpp; nullptr for unknown ids

}  // namespace psd
