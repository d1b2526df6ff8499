// Host build of sponge_b200/csrc/fp256.cuh with the C carry-chain model (PSD_HOST_EMU), exported
// for ctypes so the limb plumbing of the CUDA arithmetic can be checked without a GPU.
// Test infrastructure only.
#define PSD_HOST_EMU 1
#include "../../sponge_b200/csrc/fp256.cuh"
#include <stddef.h>

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
