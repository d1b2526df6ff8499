/* CPU oracle (C, 4 x 64-bit Montgomery limbs, OpenMP) for ark-sponge's Poseidon sponge.
 * TEST INFRASTRUCTURE ONLY -- not part of the product.  Only tests/, __graft_entry__.smoke() and
 * bench.py's cpu_baseline / --impl reference legs may load this library; sponge_b200 never does.
 *
 * It restates the reference's CPU path with the same algorithm and operation order:
 *   permutation         /root/reference/src/poseidon/mod.rs:62-118 (ark -> S-box -> dense MDS, every round)
 *   absorb / squeeze    mod.rs:121-182, 232-270, 320-342
 * on the arithmetic ark-ff performs for Fp<MontBackend<_,4>,4> without its `asm` feature
 * (Cargo.toml:22 `default-features = false`): 4-limb CIOS Montgomery multiplication on u64 limbs with
 * 128-bit products, values always fully reduced.  `pow` follows ark-ff's MSB-first
 * square-and-multiply starting from one (so alpha = 5 costs 3 squarings + 2 multiplications, as on
 * the reference's CPU path).  It is allocation-free, so as a timed baseline it is an upper bound on
 * what the reference itself (one Vec per round, parameters cloned per sponge) would reach.
 *
 * Parity status: gated in tests/ against oracle/poseidon_ref.py, which reproduces all 33
 * known-answer values of the reference's own tests (BLS12-381 Fr).  BN254 Fr and raw Montgomery
 * limbs are "parity unpinned" by the reference; they share this exact code path.
 *
 * The Rust reference cannot be compiled in this environment (no rustc/cargo, ark-ff not vendored),
 * so there is no oracle/_ref build; this port is the CPU baseline ("kind": "port").
 */
#include <stddef.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

typedef unsigned __int128 u128;

typedef struct {
    uint64_t p[4];
    uint64_t np0;    /* -p^-1 mod 2^64 */
    uint64_t one[4]; /* R mod p */
    uint64_t r2[4];  /* R^2 mod p */
    int bits;
} field_t;

typedef struct orc_ctx {
    field_t f;
    uint32_t full_rounds, partial_rounds, rate, capacity, t;
    uint64_t alpha;
    uint64_t* ark; /* Montgomery, [rounds][t][4] */
    uint64_t* mds; /* Montgomery, [t][t][4] */
} orc_ctx;

#define ORC_MAX_T 16

static int ge(const uint64_t* a, const uint64_t* b) {
    for (int i = 3; i >= 0; i--)
        if (a[i] != b[i]) return a[i] > b[i];
    return 1;
}
static void sub_p(uint64_t* r, const uint64_t* p) {
    uint64_t borrow = 0;
    for (int i = 0; i < 4; i++) {
        u128 d = (u128)r[i] - p[i] - borrow;
        r[i] = (uint64_t)d;
        borrow = (uint64_t)(d >> 64) & 1;
    }
}
/* add_assign */
static inline void fadd(const field_t* f, uint64_t* r, const uint64_t* a, const uint64_t* b) {
    u128 c = 0;
    uint64_t t[4];
    for (int i = 0; i < 4; i++) {
        c += (u128)a[i] + b[i];
        t[i] = (uint64_t)c;
        c >>= 64;
    }
    if (c || ge(t, f->p)) sub_p(t, f->p);
    memcpy(r, t, 32);
}
/* mul: CIOS */
static inline void fmul(const field_t* f, uint64_t* r, const uint64_t* a, const uint64_t* b) {
    uint64_t t[6] = {0, 0, 0, 0, 0, 0};
    for (int i = 0; i < 4; i++) {
        u128 c = 0;
        for (int j = 0; j < 4; j++) {
            c += (u128)a[j] * b[i] + t[j];
            t[j] = (uint64_t)c;
            c >>= 64;
        }
        c += t[4];
        t[4] = (uint64_t)c;
        t[5] = (uint64_t)(c >> 64);
        uint64_t m = t[0] * f->np0;
        c = (u128)m * f->p[0] + t[0];
        c >>= 64;
        for (int j = 1; j < 4; j++) {
            c += (u128)m * f->p[j] + t[j];
            t[j - 1] = (uint64_t)c;
            c >>= 64;
        }
        c += t[4];
        t[3] = (uint64_t)c;
        t[4] = t[5] + (uint64_t)(c >> 64);
    }
    if (t[4] || ge(t, f->p)) sub_p(t, f->p);
    memcpy(r, t, 32);
}
/* pow(&[alpha]): MSB-first, leading zeros skipped, starts from one (ark-ff Field::pow) */
static inline void fpow(const field_t* f, uint64_t* r, const uint64_t* a, uint64_t e) {
    uint64_t acc[4];
    memcpy(acc, f->one, 32);
    int top = 63;
    while (top >= 0 && !((e >> top) & 1)) top--;
    for (int i = top; i >= 0; i--) {
        fmul(f, acc, acc, acc);
        if ((e >> i) & 1) fmul(f, acc, acc, a);
    }
    memcpy(r, acc, 32);
}

static void field_init(field_t* f, int field_id) {
    static const uint64_t bn[4] = {0x43e1f593f0000001ull, 0x2833e84879b97091ull, 0xb85045b68181585dull,
                                   0x30644e72e131a029ull};
    /* /root/reference/src/test.rs:6 */
    static const uint64_t bls[4] = {0xffffffff00000001ull, 0x53bda402fffe5bfeull, 0x3339d80809a1d805ull,
                                    0x73eda753299d7d48ull};
    memcpy(f->p, field_id == 0 ? bn : bls, 32);
    f->bits = field_id == 0 ? 254 : 255;
    uint64_t inv = 1;
    for (int i = 0; i < 6; i++) inv *= 2 - f->p[0] * inv;
    f->np0 = (uint64_t)0 - inv;
    uint64_t x[4] = {1, 0, 0, 0};
    for (int i = 0; i < 512; i++) {
        u128 c = 0;
        for (int k = 0; k < 4; k++) {
            c += (u128)x[k] + x[k];
            x[k] = (uint64_t)c;
            c >>= 64;
        }
        if (c || ge(x, f->p)) sub_p(x, f->p);
        if (i == 255) memcpy(f->one, x, 32);
    }
    memcpy(f->r2, x, 32);
}

/* ---- context ------------------------------------------------------------------------------- */
/* ark/mds are canonical, [rounds][t][4] and [t][t][4] (PoseidonConfig::new order of meaning,
 * mod.rs:187-213).  Returns NULL on a shape the reference would reject. */
orc_ctx* orc_create(int field_id, uint32_t full_rounds, uint32_t partial_rounds, uint64_t alpha, uint32_t rate,
                    uint32_t capacity, const uint64_t* ark, const uint64_t* mds) {
    if (field_id < 0 || field_id > 1 || rate == 0 || rate + capacity > ORC_MAX_T) return NULL;
    orc_ctx* c = (orc_ctx*)calloc(1, sizeof *c);
    field_init(&c->f, field_id);
    c->full_rounds = full_rounds;
    c->partial_rounds = partial_rounds;
    c->alpha = alpha;
    c->rate = rate;
    c->capacity = capacity;
    c->t = rate + capacity;
    size_t n_ark = (size_t)(full_rounds + partial_rounds) * c->t, n_mds = (size_t)c->t * c->t;
    c->ark = (uint64_t*)malloc(n_ark * 32);
    c->mds = (uint64_t*)malloc(n_mds * 32);
    for (size_t i = 0; i < n_ark; i++) fmul(&c->f, c->ark + 4 * i, ark + 4 * i, c->f.r2);
    for (size_t i = 0; i < n_mds; i++) fmul(&c->f, c->mds + 4 * i, mds + 4 * i, c->f.r2);
    return c;
}
void orc_destroy(orc_ctx* c) {
    if (!c) return;
    free(c->ark);
    free(c->mds);
    free(c);
}

void orc_to_mont(const orc_ctx* c, uint64_t* x, size_t n) {
    for (size_t i = 0; i < n; i++) fmul(&c->f, x + 4 * i, x + 4 * i, c->f.r2);
}
void orc_from_mont(const orc_ctx* c, uint64_t* x, size_t n) {
    static const uint64_t one[4] = {1, 0, 0, 0};
    for (size_t i = 0; i < n; i++) fmul(&c->f, x + 4 * i, x + 4 * i, one);
}

/* ---- permutation, mod.rs:95-118 ------------------------------------------------------------- */
static void permute(const orc_ctx* c, uint64_t* st /* [t][4] Montgomery */) {
    const field_t* f = &c->f;
    const uint32_t t = c->t, half = c->full_rounds / 2, total = c->full_rounds + c->partial_rounds;
    uint64_t nw[ORC_MAX_T * 4];
    for (uint32_t r = 0; r < total; r++) {
        const int full = r < half || r >= half + c->partial_rounds;
        /* apply_ark 76-80 */
        for (uint32_t i = 0; i < t; i++) fadd(f, st + 4 * i, st + 4 * i, c->ark + ((size_t)r * t + i) * 4);
        /* apply_s_box 63-74 */
        if (full) {
            for (uint32_t i = 0; i < t; i++) fpow(f, st + 4 * i, st + 4 * i, c->alpha);
        } else {
            fpow(f, st, st, c->alpha);
        }
        /* apply_mds 82-93 */
        for (uint32_t i = 0; i < t; i++) {
            uint64_t cur[4] = {0, 0, 0, 0}, term[4];
            for (uint32_t j = 0; j < t; j++) {
                fmul(f, term, st + 4 * j, c->mds + ((size_t)i * t + j) * 4);
                fadd(f, cur, cur, term);
            }
            memcpy(nw + 4 * i, cur, 32);
        }
        memcpy(st, nw, (size_t)t * 32);
    }
}

/* ---- duplex sponge, mod.rs:121-182, 232-254, 320-342 ----------------------------------------- */
typedef struct {
    uint64_t st[ORC_MAX_T * 4];
    int mode; /* 0 absorbing, 1 squeezing */
    uint32_t idx;
} sponge_t;

static void sponge_new(const orc_ctx* c, sponge_t* s) {
    memset(s->st, 0, (size_t)c->t * 32);
    s->mode = 0;
    s->idx = 0;
}
static void sponge_absorb(const orc_ctx* c, sponge_t* s, const uint64_t* e /* Montgomery */, size_t n) {
    if (n == 0) return;
    uint32_t idx;
    if (s->mode == 0) {
        idx = s->idx;
        if (idx == c->rate) {
            permute(c, s->st);
            idx = 0;
        }
    } else {
        permute(c, s->st);
        idx = 0;
    }
    for (;;) {
        if (idx + n <= c->rate) {
            for (size_t i = 0; i < n; i++) {
                uint64_t* lane = s->st + 4 * (c->capacity + idx + i);
                fadd(&c->f, lane, lane, e + 4 * i);
            }
            s->mode = 0;
            s->idx = idx + (uint32_t)n;
            return;
        }
        uint32_t m = c->rate - idx;
        for (uint32_t i = 0; i < m; i++) {
            uint64_t* lane = s->st + 4 * (c->capacity + idx + i);
            fadd(&c->f, lane, lane, e + 4 * i);
        }
        permute(c, s->st);
        e += 4 * (size_t)m;
        n -= m;
        idx = 0;
    }
}
static void sponge_squeeze(const orc_ctx* c, sponge_t* s, uint64_t* out /* Montgomery */, size_t n) {
    uint32_t idx;
    if (s->mode == 0) {
        permute(c, s->st);
        idx = 0;
    } else {
        idx = s->idx;
        if (idx == c->rate) {
            permute(c, s->st);
            idx = 0;
        }
    }
    for (;;) {
        if (idx + n <= c->rate) {
            memcpy(out, s->st + 4 * (c->capacity + idx), n * 32);
            s->mode = 1;
            s->idx = idx + (uint32_t)n;
            return;
        }
        uint32_t m = c->rate - idx;
        memcpy(out, s->st + 4 * (c->capacity + idx), (size_t)m * 32);
        if (n != c->rate) permute(c, s->st); /* mod.rs:175, tested before the decrement */
        out += 4 * (size_t)m;
        n -= m;
        idx = 0;
    }
}

/* ---- batch entry points (AoS, Montgomery unless stated) --------------------------------------- */
static int pick_threads(int threads) {
#ifdef _OPENMP
    return threads > 0 ? threads : omp_get_max_threads();
#else
    (void)threads;
    return 1;
#endif
}
int orc_max_threads(void) { return pick_threads(0); }

void orc_permute_batch(const orc_ctx* c, uint64_t* states, size_t n, int threads) {
    const int nt = pick_threads(threads);
    (void)nt;
#pragma omp parallel for num_threads(nt) schedule(static)
    for (long long i = 0; i < (long long)n; i++) permute(c, states + (size_t)i * c->t * 4);
}

void orc_absorb_squeeze_batch(const orc_ctx* c, const uint64_t* in, size_t n, uint32_t n_absorb, uint32_t n_squeeze,
                              uint64_t* out, int threads) {
    const int nt = pick_threads(threads);
    (void)nt;
#pragma omp parallel for num_threads(nt) schedule(static)
    for (long long i = 0; i < (long long)n; i++) {
        sponge_t s;
        sponge_new(c, &s);
        sponge_absorb(c, &s, in + (size_t)i * n_absorb * 4, n_absorb);
        sponge_squeeze(c, &s, out + (size_t)i * n_squeeze * 4, n_squeeze);
    }
}

/* squeeze_bytes, mod.rs:256-270 */
void orc_absorb_squeeze_bytes_batch(const orc_ctx* c, const uint64_t* in, size_t n, uint32_t n_absorb, uint32_t n_bytes,
                                    uint8_t* out, int threads) {
    const int nt = pick_threads(threads);
    (void)nt;
    const uint32_t usable = (uint32_t)(c->f.bits - 1) / 8;
    const uint32_t n_el = (n_bytes + usable - 1) / usable;
    static const uint64_t one[4] = {1, 0, 0, 0};
#pragma omp parallel for num_threads(nt) schedule(static)
    for (long long i = 0; i < (long long)n; i++) {
        sponge_t s;
        sponge_new(c, &s);
        sponge_absorb(c, &s, in + (size_t)i * n_absorb * 4, n_absorb);
        uint64_t* el = (uint64_t*)malloc((size_t)(n_el ? n_el : 1) * 32);
        sponge_squeeze(c, &s, el, n_el);
        for (uint32_t e = 0; e < n_el; e++) {
            uint64_t canon[4];
            fmul(&c->f, canon, el + 4 * e, one);
            for (uint32_t b = 0; b < usable && e * usable + b < n_bytes; b++)
                out[(size_t)i * n_bytes + e * usable + b] = (uint8_t)(canon[b / 8] >> (8 * (b % 8)));
        }
        free(el);
    }
}

/* H(l,r) = new; absorb([l,r]); squeeze_native(1)[0]; level by level */
void orc_merkle_root(const orc_ctx* c, const uint64_t* leaves, size_t n_leaves, uint64_t* root, int threads) {
    if (n_leaves == 1) {
        memcpy(root, leaves, 32);
        return;
    }
    uint64_t* cur = (uint64_t*)malloc(n_leaves / 2 * 32);
    orc_absorb_squeeze_batch(c, leaves, n_leaves / 2, 2, 1, cur, threads);
    for (size_t n = n_leaves / 2; n > 1; n /= 2) {
        uint64_t* nxt = (uint64_t*)malloc(n / 2 * 32);
        orc_absorb_squeeze_batch(c, cur, n / 2, 2, 1, nxt, threads);
        free(cur);
        cur = nxt;
    }
    memcpy(root, cur, 32);
    free(cur);
}

/* ---- a single sponge driven call by call (state-machine fuzzing) ------------------------------ */
typedef struct orc_sponge {
    const orc_ctx* c;
    sponge_t s;
} orc_sponge;

orc_sponge* orc_sponge_new(const orc_ctx* c) {
    orc_sponge* s = (orc_sponge*)calloc(1, sizeof *s);
    s->c = c;
    sponge_new(c, &s->s);
    return s;
}
void orc_sponge_free(orc_sponge* s) { free(s); }
void orc_sponge_absorb(orc_sponge* s, const uint64_t* e, size_t n) { sponge_absorb(s->c, &s->s, e, n); }
void orc_sponge_squeeze(orc_sponge* s, uint64_t* out, size_t n) { sponge_squeeze(s->c, &s->s, out, n); }
void orc_sponge_state(const orc_sponge* s, uint64_t* st, int* mode, uint32_t* idx) {
    memcpy(st, s->s.st, (size_t)s->c->t * 32);
    *mode = s->s.mode;
    *idx = s->s.idx;
}

/* ---- synthetic inputs; must match poseidon_ref.py::synth_element -------------- */
static uint64_t splitmix64(uint64_t x) {
    x += 0x9E3779B97F4A7C15ull;
    uint64_t z = x;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}
/* writes n elements starting at index `first`; montgomery != 0 -> Montgomery form */
void orc_synth_fill(const orc_ctx* c, uint64_t* out, size_t n, uint64_t seed, uint64_t first, int montgomery,
                    int threads) {
    const int nt = pick_threads(threads);
    (void)nt;
    const uint64_t top_mask = (1ull << (c->f.bits - 192)) - 1;
#pragma omp parallel for num_threads(nt) schedule(static)
    for (long long i = 0; i < (long long)n; i++) {
        uint64_t* e = out + 4 * (size_t)i;
        const uint64_t k = first + (uint64_t)i;
        for (uint64_t salt = 0;; salt++) {
            for (int w = 0; w < 4; w++) e[w] = splitmix64(((seed << 40) ^ (salt << 34)) + 4 * k + (uint64_t)w);
            e[3] &= top_mask;
            if (!ge(e, c->f.p)) break;
        }
        if (montgomery) fmul(&c->f, e, e, c->f.r2);
    }
}
