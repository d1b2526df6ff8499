// Exercises include/poseidon_b200.hpp (the C++ mirror of CryptographicSponge) against the reference's
// own known answer (src/poseidon/mod.rs:376-399).  argv[1..3] = expected canonical outputs as 64 hex digits.
// Exit codes: 0 ok, 3 = no CUDA device (library refused to run: it has no CPU path), else failure.
#include <cstdio>
#include <cstring>
#include <string>

#include "poseidon_b200.hpp"

static psd::Fe from_hex(const char* h) {
    psd::Fe f;
    std::string s(h);
    for (int l = 0; l < 4; l++) f.limbs[l] = std::stoull(s.substr(64 - 16 * (l + 1), 16), nullptr, 16);
    return f;
}

int main(int argc, char** argv) {
    using F = psd::Bls12_381Fr;
    auto cfg = psd::PoseidonConfig<F>::get_default_poseidon_parameters(2, false);
    if (!cfg || cfg->alpha != 17 || cfg->partial_rounds != 31) return 1;
    if (psd::PoseidonConfig<F>::get_default_poseidon_parameters(9, false)) return 1;   // None in the reference
    try {
        auto ctx = std::make_shared<psd::Context<F>>(*cfg);
        psd::PoseidonSponge<F> sponge(ctx);
        sponge.absorb({psd::Fe::from_u64(0), psd::Fe::from_u64(1), psd::Fe::from_u64(2)});
        auto clone = sponge.clone();
        auto res = sponge.squeeze_native_field_elements(3);
        for (int i = 0; i < 3 && i + 1 < argc; i++)
            if (!(res[i] == from_hex(argv[i + 1]))) { std::printf("KAT mismatch at %d\n", i); return 2; }
        // squeeze_field_elements::<F> on a clone gives the same values (tests.rs:71-85)
        if (!(clone.squeeze_field_elements<F>(3) == res)) return 2;
        auto st = sponge.into_state();
        auto again = psd::PoseidonSponge<F>::from_state(st, ctx);
        if (again.squeeze_bytes(40) != sponge.squeeze_bytes(40)) return 2;
        auto forked = sponge.fork({1, 2, 3});
        if (forked.squeeze_bits(100) == sponge.squeeze_bits(100)) return 2;
        psd::Fe states[6] = {};
        states[4] = psd::Fe::from_u64(1);
        ctx->permute_batch(states, 2, PSD_REPR_CANONICAL);
        std::printf("ok\n");
    } catch (const psd::Error& e) {
        std::printf("%s\n", e.what());
        return e.status == PSD_ERR_CUDA ? 3 : 2;
    }
    return 0;
}
