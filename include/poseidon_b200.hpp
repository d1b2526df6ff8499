// poseidon_b200.hpp -- C++17 RAII mirror of ark-sponge's Poseidon interface over the C ABI
// (include/poseidon_b200.h).  Header-only; link with -lposeidon_b200.
//
// Same names and argument meaning as the reference (paths relative to /root/reference/src):
//   PoseidonConfig<F>            poseidon/mod.rs:23-42, ::new 187-213 (mds BEFORE ark)
//   PoseidonSponge<F>            poseidon/mod.rs:44-60; CryptographicSponge lib.rs:105-158 (impl mod.rs:216-318);
//                                FieldBasedCryptographicSponge lib.rs:162-183 (impl mod.rs:320-342);
//                                SpongeExt lib.rs:188-195 (impl mod.rs:344-367)
//   DuplexSpongeMode             lib.rs:198-210
// plus the batch entry points (permute_batch / absorb_squeeze_batch / merkle_root).  Errors: the
// reference panics; here every failing call throws psd::Error carrying the psd_status.
#pragma once
#include <algorithm>
#include <array>
#include <cstdint>
#include <memory>
#include <stdexcept>
#include <string>
#include <vector>

#include "poseidon_b200.h"

namespace psd {

struct Error : std::runtime_error {
    psd_status status;
    Error(psd_status s, const std::string& m) : std::runtime_error(std::string(psd_status_string(s)) + ": " + m), status(s) {}
};
inline void check(psd_status s) {
    if (s != PSD_OK) throw Error(s, psd_last_error_string());
}

// One field element: the memory image of ark_ff::Fp<MontBackend<_,4>,4> (4 LE u64 limbs).
struct Fe {
    std::array<uint64_t, 4> limbs{};
    static Fe from_u64(uint64_t v) { Fe f; f.limbs[0] = v; return f; }   // canonical small value
    bool operator==(const Fe& o) const { return limbs == o.limbs; }
};
static_assert(sizeof(Fe) == 32, "Fe must be 32 bytes");

// Field tags play the role of the reference's type parameter F.
struct Bn254Fr { static constexpr psd_field id = PSD_FIELD_BN254_FR; static constexpr unsigned modulus_bit_size = 254; };
struct Bls12_381Fr { static constexpr psd_field id = PSD_FIELD_BLS12_381_FR; static constexpr unsigned modulus_bit_size = 255; };

// lib.rs:198-210
struct DuplexSpongeMode {
    enum Kind { Absorbing, Squeezing } kind = Absorbing;
    uint32_t index = 0;   // next_absorb_index / next_squeeze_index
};

// poseidon/mod.rs:23-42.  ark[round][lane], mds[row][col], canonical values.
template <class F>
struct PoseidonConfig {
    size_t full_rounds = 0, partial_rounds = 0;
    uint64_t alpha = 0;
    std::vector<std::vector<Fe>> ark, mds;
    size_t rate = 0, capacity = 0;

    // PoseidonConfig::new, mod.rs:187-213: note the argument order, mds before ark.
    static PoseidonConfig create(size_t full_rounds, size_t partial_rounds, uint64_t alpha,
                                 std::vector<std::vector<Fe>> mds, std::vector<std::vector<Fe>> ark, size_t rate,
                                 size_t capacity) {
        if (ark.size() != full_rounds + partial_rounds) throw Error(PSD_ERR_SHAPE, "ark.len() != full_rounds + partial_rounds");
        for (auto& r : ark) if (r.size() != rate + capacity) throw Error(PSD_ERR_SHAPE, "ark row length");
        if (mds.size() != rate + capacity) throw Error(PSD_ERR_SHAPE, "mds.len()");
        for (auto& r : mds) if (r.size() != rate + capacity) throw Error(PSD_ERR_SHAPE, "mds row length");
        PoseidonConfig c;
        c.full_rounds = full_rounds; c.partial_rounds = partial_rounds; c.alpha = alpha;
        c.mds = std::move(mds); c.ark = std::move(ark); c.rate = rate; c.capacity = capacity;
        return c;
    }

    // PoseidonDefaultConfigField::get_default_poseidon_parameters, traits.rs:59-102; empty optional -> nullptr
    static std::unique_ptr<PoseidonConfig> get_default_poseidon_parameters(size_t rate, bool optimized_for_weights) {
        std::vector<uint64_t> ark(73 * 9 * 4), mds(81 * 4);
        psd_params p{};
        psd_status st = psd_default_params(F::id, (uint32_t)rate, optimized_for_weights, &p, ark.data(), mds.data());
        if (st == PSD_ERR_UNSUPPORTED) return nullptr;
        check(st);
        auto c = std::make_unique<PoseidonConfig>();
        const size_t t = p.rate + p.capacity, rounds = p.full_rounds + p.partial_rounds;
        c->full_rounds = p.full_rounds; c->partial_rounds = p.partial_rounds; c->alpha = p.alpha;
        c->rate = p.rate; c->capacity = p.capacity;
        c->ark.assign(rounds, std::vector<Fe>(t));
        c->mds.assign(t, std::vector<Fe>(t));
        for (size_t i = 0; i < rounds * t; i++) std::copy_n(&ark[4 * i], 4, c->ark[i / t][i % t].limbs.begin());
        for (size_t i = 0; i < t * t; i++) std::copy_n(&mds[4 * i], 4, c->mds[i / t][i % t].limbs.begin());
        return c;
    }
};
template <class F> using PoseidonParameters = PoseidonConfig<F>;   // ark-sponge 0.3.0 name

// psd_ctx: the parameter set resident on one GPU (replaces parameters.clone() per sponge, mod.rs:226)
template <class F>
class Context {
   public:
    explicit Context(const PoseidonConfig<F>& cfg, int device = 0) : t_(cfg.rate + cfg.capacity) {
        std::vector<Fe> ark, mds;
        for (auto& r : cfg.ark) ark.insert(ark.end(), r.begin(), r.end());
        for (auto& r : cfg.mds) mds.insert(mds.end(), r.begin(), r.end());
        psd_params p{};
        p.field = F::id; p.full_rounds = (uint32_t)cfg.full_rounds; p.partial_rounds = (uint32_t)cfg.partial_rounds;
        p.alpha = cfg.alpha; p.rate = (uint32_t)cfg.rate; p.capacity = (uint32_t)cfg.capacity;
        p.ark = reinterpret_cast<const uint64_t*>(ark.data()); p.mds = reinterpret_cast<const uint64_t*>(mds.data());
        p.repr = PSD_REPR_CANONICAL;
        psd_ctx* h = nullptr;
        check(psd_ctx_create(&p, device, &h));
        h_.reset(h, psd_ctx_destroy);
    }
    psd_ctx* get() const { return h_.get(); }
    size_t width() const { return t_; }

    // batch entry points over host memory (AoS); repr as given
    void permute_batch(Fe* states, size_t n, psd_repr repr = PSD_REPR_MONTGOMERY) const {
        check(psd_permute_batch(get(), reinterpret_cast<uint64_t*>(states), n, PSD_LAYOUT_AOS, repr, PSD_MEM_HOST, nullptr));
    }
    std::vector<Fe> absorb_squeeze_batch(const Fe* inputs, size_t n, uint32_t n_absorb, uint32_t n_squeeze,
                                         psd_repr repr = PSD_REPR_MONTGOMERY) const {
        std::vector<Fe> out(n * n_squeeze);
        check(psd_absorb_squeeze_batch(get(), reinterpret_cast<const uint64_t*>(inputs), n, n_absorb, n_squeeze,
                                       reinterpret_cast<uint64_t*>(out.data()), PSD_LAYOUT_AOS, repr, PSD_MEM_HOST, nullptr));
        return out;
    }
    Fe merkle_root(const Fe* leaves, size_t n_leaves, psd_repr repr = PSD_REPR_MONTGOMERY) const {
        Fe root;
        check(psd_merkle_root(get(), reinterpret_cast<const uint64_t*>(leaves), n_leaves, root.limbs.data(), repr,
                              PSD_MEM_HOST, nullptr));
        return root;
    }

   private:
    std::shared_ptr<psd_ctx> h_;
    size_t t_;
};

// PoseidonSponge<F>: values are canonical Fe on this surface.
template <class F>
class PoseidonSponge {
   public:
    // CryptographicSponge::new, mod.rs:219-230
    explicit PoseidonSponge(std::shared_ptr<Context<F>> ctx) : ctx_(std::move(ctx)) {
        psd_sponge* s = nullptr;
        check(psd_sponge_new(ctx_->get(), &s));
        s_.reset(s);
    }
    // Clone, mod.rs:44
    PoseidonSponge clone() const {
        psd_sponge* s = nullptr;
        check(psd_sponge_clone(s_.get(), &s));
        return PoseidonSponge(ctx_, s);
    }
    // absorb(&[F]) / absorb(&F), mod.rs:232-254
    void absorb(const std::vector<Fe>& elems) { check(psd_sponge_absorb(s_.get(), reinterpret_cast<const uint64_t*>(elems.data()), elems.size(), PSD_REPR_CANONICAL)); }
    void absorb(const Fe& e) { check(psd_sponge_absorb(s_.get(), e.limbs.data(), 1, PSD_REPR_CANONICAL)); }
    // absorb(&[u8]), absorb.rs:122-140
    void absorb_bytes(const std::vector<uint8_t>& bytes) { check(psd_sponge_absorb_bytes(s_.get(), bytes.data(), bytes.size())); }
    // squeeze_native_field_elements, mod.rs:320-342
    std::vector<Fe> squeeze_native_field_elements(size_t n) {
        std::vector<Fe> out(n ? n : 1);
        check(psd_sponge_squeeze_native(s_.get(), n, reinterpret_cast<uint64_t*>(out.data()), PSD_REPR_CANONICAL));
        out.resize(n);
        return out;
    }
    // squeeze_bytes, mod.rs:256-270
    std::vector<uint8_t> squeeze_bytes(size_t n) {
        std::vector<uint8_t> out(n ? n : 1);
        check(psd_sponge_squeeze_bytes(s_.get(), n, out.data()));
        out.resize(n);
        return out;
    }
    // squeeze_bits, mod.rs:272-286
    std::vector<bool> squeeze_bits(size_t n) {
        std::vector<uint8_t> raw(n ? n : 1);
        check(psd_sponge_squeeze_bits(s_.get(), n, raw.data()));
        return std::vector<bool>(raw.begin(), raw.begin() + n);
    }
    // squeeze_field_elements::<F2>, mod.rs:306-317
    template <class F2>
    std::vector<Fe> squeeze_field_elements(size_t n) {
        std::vector<Fe> out(n ? n : 1);
        check(psd_sponge_squeeze_field_elements(s_.get(), F2::id, n, reinterpret_cast<uint64_t*>(out.data())));
        out.resize(n);
        return out;
    }
    // fork, lib.rs:149-157
    PoseidonSponge fork(const std::vector<uint8_t>& domain) const {
        psd_sponge* s = nullptr;
        check(psd_sponge_fork(s_.get(), domain.data(), domain.size(), &s));
        return PoseidonSponge(ctx_, s);
    }
    // SpongeExt, lib.rs:188-195
    struct State { std::vector<Fe> state; DuplexSpongeMode mode; };
    State into_state() const {
        State st;
        st.state.resize(ctx_->width());
        int mode = 0; uint32_t idx = 0;
        check(psd_sponge_get_state(s_.get(), reinterpret_cast<uint64_t*>(st.state.data()), PSD_REPR_CANONICAL, &mode, &idx));
        st.mode.kind = mode == PSD_MODE_ABSORBING ? DuplexSpongeMode::Absorbing : DuplexSpongeMode::Squeezing;
        st.mode.index = idx;
        return st;
    }
    static PoseidonSponge from_state(const State& st, std::shared_ptr<Context<F>> ctx) {
        PoseidonSponge s(std::move(ctx));
        check(psd_sponge_set_state(s.s_.get(), reinterpret_cast<const uint64_t*>(st.state.data()), PSD_REPR_CANONICAL,
                                   st.mode.kind == DuplexSpongeMode::Absorbing ? PSD_MODE_ABSORBING : PSD_MODE_SQUEEZING,
                                   st.mode.index));
        return s;
    }

   private:
    struct Del { void operator()(psd_sponge* p) const { psd_sponge_free(p); } };
    PoseidonSponge(std::shared_ptr<Context<F>> ctx, psd_sponge* s) : ctx_(std::move(ctx)), s_(s) {}
    std::shared_ptr<Context<F>> ctx_;
    std::unique_ptr<psd_sponge, Del> s_;
};

}  // namespace psd
