//! The batch entry points ark-sponge does not have: N independent sponges per call.
//! NOT COMPILED here (see lib.rs).
use crate::{check, ffi, Ctx};
use ark_ff::{Fp, FpConfig, Zero};
use ark_sponge::poseidon::PoseidonConfig;
use std::sync::Arc;

/// N independent `PoseidonSponge::permute`s (poseidon/mod.rs:95-118) over a slice of states, in place.
pub fn permute_batch<P: FpConfig<4>>(ctx: &Arc<Ctx>, cfg: &PoseidonConfig<Fp<P, 4>>, states: &mut [Fp<P, 4>]) {
    let t = cfg.rate + cfg.capacity;
    assert_eq!(states.len() % t, 0, "states must hold a whole number of {}-element states", t);
    check(unsafe {
        ffi::psd_permute_batch(ctx.raw(), states.as_mut_ptr() as *mut u64, states.len() / t, ffi::LAYOUT_AOS,
                               ffi::REPR_MONTGOMERY, ffi::MEM_HOST, std::ptr::null_mut())
    });
}

/// For every chunk of `n_absorb` inputs: `new(); absorb(chunk); squeeze_native_field_elements(n_squeeze)`.
/// Replaces `inputs.par_chunks(k).map(|c| { let mut s = PoseidonSponge::new(&cfg); s.absorb(&c); s.squeeze_native_field_elements(m) })`.
pub fn absorb_squeeze_batch<P: FpConfig<4>>(ctx: &Arc<Ctx>, inputs: &[Fp<P, 4>], n_absorb: usize, n_squeeze: usize) -> Vec<Fp<P, 4>> {
    assert!(n_absorb > 0 && inputs.len() % n_absorb == 0);
    let n = inputs.len() / n_absorb;
    let mut out = vec![Fp::<P, 4>::zero(); n * n_squeeze];
    check(unsafe {
        ffi::psd_absorb_squeeze_batch(ctx.raw(), inputs.as_ptr() as *const u64, n, n_absorb as u32, n_squeeze as u32,
                                      out.as_mut_ptr() as *mut u64, ffi::LAYOUT_AOS, ffi::REPR_MONTGOMERY, ffi::MEM_HOST,
                                      std::ptr::null_mut())
    });
    out
}

/// Merkle root with `H(l, r) = { new(); absorb(&[l, r]); squeeze_native_field_elements(1)[0] }`; `leaves.len()` a power of two.
pub fn merkle_root<P: FpConfig<4>>(ctx: &Arc<Ctx>, leaves: &[Fp<P, 4>]) -> Fp<P, 4> {
    let mut root = Fp::<P, 4>::zero();
    check(unsafe {
        ffi::psd_merkle_root(ctx.raw(), leaves.as_ptr() as *const u64, leaves.len(), &mut root as *mut Fp<P, 4> as *mut u64,
                             ffi::REPR_MONTGOMERY, ffi::MEM_HOST, std::ptr::null_mut())
    });
    root
}
