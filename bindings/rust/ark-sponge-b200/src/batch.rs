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

/// Sponge i absorbs `inputs[i]` (lists of differing lengths, as in the reference's own behavioural test,
/// poseidon/tests.rs:45-69) and squeezes `n_squeeze` native elements.
pub fn absorb_squeeze_ragged<P: FpConfig<4>>(ctx: &Arc<Ctx>, inputs: &[Vec<Fp<P, 4>>], n_squeeze: usize) -> Vec<Fp<P, 4>> {
    let mut offsets: Vec<u64> = Vec::with_capacity(inputs.len() + 1);
    let mut flat: Vec<Fp<P, 4>> = Vec::new();
    offsets.push(0);
    for v in inputs {
        flat.extend_from_slice(v);
        offsets.push(flat.len() as u64);
    }
    let mut out = vec![Fp::<P, 4>::zero(); inputs.len() * n_squeeze];
    check(unsafe {
        ffi::psd_absorb_squeeze_ragged_batch(ctx.raw(), flat.as_ptr() as *const u64, offsets.as_ptr(), inputs.len(), n_squeeze as u32,
                                             out.as_mut_ptr() as *mut u64, ffi::REPR_MONTGOMERY, ffi::MEM_HOST, std::ptr::null_mut())
    });
    out
}

/// `permute_batch` over every GPU of the box: one context per device (cached), contiguous shards, no exchange.
pub fn permute_batch_all_gpus<P: FpConfig<4>>(cfg: &PoseidonConfig<Fp<P, 4>>, states: &mut [Fp<P, 4>]) {
    let t = cfg.rate + cfg.capacity;
    assert_eq!(states.len() % t, 0);
    let n_gpus = unsafe { ffi::psd_device_count() };
    let ctxs: Vec<Arc<Ctx>> = (0..n_gpus).map(|d| Ctx::cached(cfg, d)).collect();
    let raw: Vec<*mut ffi::psd_ctx> = ctxs.iter().map(|c| c.raw()).collect();
    check(unsafe {
        ffi::psd_permute_batch_multi(raw.as_ptr(), n_gpus, states.as_mut_ptr() as *mut u64, states.len() / t, ffi::LAYOUT_AOS,
                                     ffi::REPR_MONTGOMERY)
    });
}

/// One Merkle tree over every GPU of the box: per-GPU subtrees, the subtree roots gathered on the first device.
/// `leaves.len()` and the device count must be powers of two.
pub fn merkle_root_all_gpus<P: FpConfig<4>>(cfg: &PoseidonConfig<Fp<P, 4>>, leaves: &[Fp<P, 4>]) -> Fp<P, 4> {
    let n_gpus = unsafe { ffi::psd_device_count() } as usize;
    let per = leaves.len() / n_gpus;
    let ctxs: Vec<Arc<Ctx>> = (0..n_gpus).map(|d| Ctx::cached(cfg, d as i32)).collect();
    let raw: Vec<*mut ffi::psd_ctx> = ctxs.iter().map(|c| c.raw()).collect();
    let ptrs: Vec<*const u64> = (0..n_gpus).map(|g| leaves[g * per..].as_ptr() as *const u64).collect();
    let mut root = Fp::<P, 4>::zero();
    check(unsafe {
        ffi::psd_merkle_root_multi(raw.as_ptr(), n_gpus as i32, ptrs.as_ptr(), per, &mut root as *mut Fp<P, 4> as *mut u64,
                                   ffi::REPR_MONTGOMERY, ffi::MEM_HOST, std::ptr::null())
    });
    root
}
