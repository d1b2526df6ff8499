//! Raw declarations of include/poseidon_b200.h (only the entry points this crate uses).
#![allow(non_camel_case_types, dead_code)]
use std::os::raw::{c_char, c_int, c_void};

#[repr(C)]
pub struct psd_ctx {
    _private: [u8; 0],
}
#[repr(C)]
pub struct psd_sponge {
    _private: [u8; 0],
}

/// Mirrors `PoseidonConfig<F>` (ark-sponge src/poseidon/mod.rs:23-42).
#[repr(C)]
pub struct psd_params {
    pub field: i32,
    pub full_rounds: u32,
    pub partial_rounds: u32,
    pub alpha: u64,
    pub rate: u32,
    pub capacity: u32,
    pub ark: *const u64,
    pub mds: *const u64,
    pub repr: i32,
}

pub const PSD_OK: c_int = 0;
pub const FIELD_BN254_FR: i32 = 0;
pub const FIELD_BLS12_381_FR: i32 = 1;
pub const REPR_CANONICAL: c_int = 0;
pub const REPR_MONTGOMERY: c_int = 1;
pub const LAYOUT_AOS: c_int = 0;
pub const MEM_HOST: c_int = 0;
pub const MODE_ABSORBING: c_int = 0;
pub const MODE_SQUEEZING: c_int = 1;

extern "C" {
    pub fn psd_last_error_string() -> *const c_char;
    pub fn psd_ctx_create(params: *const psd_params, device: c_int, out: *mut *mut psd_ctx) -> c_int;
    pub fn psd_ctx_destroy(ctx: *mut psd_ctx);
    pub fn psd_sponge_new(ctx: *mut psd_ctx, out: *mut *mut psd_sponge) -> c_int;
    pub fn psd_sponge_clone(s: *const psd_sponge, out: *mut *mut psd_sponge) -> c_int;
    pub fn psd_sponge_free(s: *mut psd_sponge);
    pub fn psd_sponge_absorb(s: *mut psd_sponge, elems: *const u64, n: usize, repr: c_int) -> c_int;
    pub fn psd_sponge_squeeze_native(s: *mut psd_sponge, n: usize, out: *mut u64, repr: c_int) -> c_int;
    pub fn psd_sponge_squeeze_bytes(s: *mut psd_sponge, n_bytes: usize, out: *mut u8) -> c_int;
    pub fn psd_sponge_squeeze_bits(s: *mut psd_sponge, n_bits: usize, out: *mut u8) -> c_int;
    pub fn psd_sponge_get_state(s: *const psd_sponge, state: *mut u64, repr: c_int, mode: *mut c_int, index: *mut u32) -> c_int;
    pub fn psd_sponge_set_state(s: *mut psd_sponge, state: *const u64, repr: c_int, mode: c_int, index: u32) -> c_int;
    pub fn psd_permute_batch(ctx: *mut psd_ctx, states: *mut u64, n: usize, layout: c_int, repr: c_int, mem: c_int, stream: *mut c_void) -> c_int;
    pub fn psd_absorb_squeeze_batch(ctx: *mut psd_ctx, inputs: *const u64, n: usize, n_absorb: u32, n_squeeze: u32, outputs: *mut u64, layout: c_int, repr: c_int, mem: c_int, stream: *mut c_void) -> c_int;
    pub fn psd_merkle_root(ctx: *mut psd_ctx, leaves: *const u64, n_leaves: usize, root: *mut u64, repr: c_int, mem: c_int, stream: *mut c_void) -> c_int;
    pub fn psd_absorb_squeeze_ragged_batch(ctx: *mut psd_ctx, inputs: *const u64, offsets: *const u64, n: usize, n_squeeze: u32, outputs: *mut u64, repr: c_int, mem: c_int, stream: *mut c_void) -> c_int;
    pub fn psd_device_count() -> c_int;
    pub fn psd_permute_batch_multi(ctxs: *const *mut psd_ctx, n_gpus: c_int, states: *mut u64, n: usize, layout: c_int, repr: c_int) -> c_int;
    pub fn psd_absorb_squeeze_batch_multi(ctxs: *const *mut psd_ctx, n_gpus: c_int, inputs: *const u64, n: usize, n_absorb: u32, n_squeeze: u32, outputs: *mut u64, layout: c_int, repr: c_int) -> c_int;
    pub fn psd_merkle_root_multi(ctxs: *const *mut psd_ctx, n_gpus: c_int, leaves_per_gpu: *const *const u64, leaves_per_gpu_n: usize, root: *mut u64, repr: c_int, mem: c_int, streams: *const *mut c_void) -> c_int;
    pub fn psd_host_alloc(bytes: usize, out: *mut *mut c_void) -> c_int;
    pub fn psd_host_free(p: *mut c_void);
}
