//! `B200PoseidonSponge<Fp<P,4>>`: ark-sponge's `CryptographicSponge`, `FieldBasedCryptographicSponge` and
//! `SpongeExt` for Poseidon, executed by libposeidon_b200 on an NVIDIA B200.
//!
//! NOT COMPILED in this repository's environment (no Rust toolchain, ark-ff not vendored) -- written
//! against the ark-ff 0.4 / ark-sponge 0.3 API the reference uses.  Each method names the reference
//! code it replaces (paths relative to ark-sponge's src/).
//!
//! An `Fp<MontBackend<_,4>,4>` is `Fp(BigInt([u64;4]), PhantomData)` holding the Montgomery residue in four
//! little-endian limbs: exactly one 32-byte element of the C ABI with PSD_REPR_MONTGOMERY, so slices of field
//! elements cross the boundary as raw pointers with no conversion.
mod ffi;
pub mod batch;

use ark_ff::{BigInteger, Fp, FpConfig, PrimeField};
use ark_sponge::poseidon::{PoseidonConfig, PoseidonSpongeState};
use ark_sponge::{Absorb, CryptographicSponge, DuplexSpongeMode, FieldBasedCryptographicSponge, FieldElementSize, SpongeExt};
use std::collections::HashMap;
use std::ffi::CStr;
use std::os::raw::c_int;
use std::sync::{Arc, Mutex, OnceLock, Weak};

/// Turns a status code back into the panic the reference would have raised (shape asserts
/// poseidon/mod.rs:196-203, unwraps absorb.rs:163); the drop-in must not change error behaviour.
pub(crate) fn check(rc: c_int) {
    if rc != ffi::PSD_OK {
        let msg = unsafe { CStr::from_ptr(ffi::psd_last_error_string()) }.to_string_lossy().into_owned();
        panic!("poseidon_b200: {}", msg);
    }
}

pub(crate) fn field_id<F: PrimeField>() -> i32 {
    const BN254: &str = "21888242871839275222246405745257275088548364400416034343698204186575808495617";
    const BLS: &str = "52435875175126190479447740508185965837690552500527637822603658699938581184513";
    let m = F::MODULUS.to_string();
    if m == BN254 {
        ffi::FIELD_BN254_FR
    } else if m == BLS {
        ffi::FIELD_BLS12_381_FR
    } else {
        panic!("poseidon_b200 supports BN254 Fr and BLS12-381 Fr only")
    }
}

/// One parameter set resident on the GPU (psd_ctx).  Replaces the per-sponge `parameters.clone()`
/// of `PoseidonSponge::new` (poseidon/mod.rs:226).
pub struct Ctx {
    raw: *mut ffi::psd_ctx,
}
unsafe impl Send for Ctx {}
unsafe impl Sync for Ctx {} // immutable after creation (see the header's threading notes)
impl Drop for Ctx {
    fn drop(&mut self) {
        unsafe { ffi::psd_ctx_destroy(self.raw) }
    }
}
/// Key of the process-wide context cache: device, field and every number of the parameter set.
#[derive(PartialEq, Eq, Hash, Clone)]
struct CtxKey {
    device: i32,
    field: i32,
    shape: [u64; 5],   // full_rounds, partial_rounds, alpha, rate, capacity
    limbs: Vec<u64>,   // ark then mds, Montgomery limb images
}

fn ctx_cache() -> &'static Mutex<HashMap<CtxKey, Weak<Ctx>>> {
    static CACHE: OnceLock<Mutex<HashMap<CtxKey, Weak<Ctx>>>> = OnceLock::new();
    CACHE.get_or_init(|| Mutex::new(HashMap::new()))
}

impl Ctx {
    /// The context of this parameter set on `device`: the live one if any sponge or caller still holds
    /// it, else a new one.  The reference calls `PoseidonSponge::new(&params)` freely and keeps many
    /// sponges alive at once (poseidon/mod.rs:219-230 clones the parameters into each); here all of them
    /// share one psd_ctx (streams, staging buffers, constants on the GPU) instead of creating one each.
    pub fn cached<P: FpConfig<4>>(p: &PoseidonConfig<Fp<P, 4>>, device: i32) -> Arc<Self> {
        let ark: Vec<Fp<P, 4>> = p.ark.iter().flatten().copied().collect();
        let mds: Vec<Fp<P, 4>> = p.mds.iter().flatten().copied().collect();
        let mut limbs: Vec<u64> = Vec::with_capacity(4 * (ark.len() + mds.len()));
        for e in ark.iter().chain(mds.iter()) {
            limbs.extend_from_slice(&(e.0).0);
        }
        let key = CtxKey {
            device,
            field: field_id::<Fp<P, 4>>(),
            shape: [p.full_rounds as u64, p.partial_rounds as u64, p.alpha, p.rate as u64, p.capacity as u64],
            limbs,
        };
        let mut map = ctx_cache().lock().unwrap();
        if let Some(live) = map.get(&key).and_then(Weak::upgrade) {
            return live;
        }
        map.retain(|_, w| w.strong_count() > 0);
        let fresh = Self::new(p, device);
        map.insert(key, Arc::downgrade(&fresh));
        fresh
    }

    /// Always creates a new psd_ctx (identical parameter sets still share one copy of the constants on
    /// the GPU inside the library).
    pub fn new<P: FpConfig<4>>(p: &PoseidonConfig<Fp<P, 4>>, device: i32) -> Arc<Self> {
        // [round][lane] and [row][col], flattened; the Vec<Fp> memory is the Montgomery limb image
        let ark: Vec<Fp<P, 4>> = p.ark.iter().flatten().copied().collect();
        let mds: Vec<Fp<P, 4>> = p.mds.iter().flatten().copied().collect();
        let params = ffi::psd_params {
            field: field_id::<Fp<P, 4>>(),
            full_rounds: p.full_rounds as u32,
            partial_rounds: p.partial_rounds as u32,
            alpha: p.alpha,
            rate: p.rate as u32,
            capacity: p.capacity as u32,
            ark: ark.as_ptr() as *const u64,
            mds: mds.as_ptr() as *const u64,
            repr: ffi::REPR_MONTGOMERY,
        };
        let mut raw = std::ptr::null_mut();
        check(unsafe { ffi::psd_ctx_create(&params, device, &mut raw) });
        Arc::new(Ctx { raw })
    }
    pub(crate) fn raw(&self) -> *mut ffi::psd_ctx {
        self.raw
    }
}

pub struct B200PoseidonSponge<P: FpConfig<4>> {
    pub parameters: PoseidonConfig<Fp<P, 4>>,
    ctx: Arc<Ctx>,
    h: *mut ffi::psd_sponge,
}

impl<P: FpConfig<4>> Drop for B200PoseidonSponge<P> {
    fn drop(&mut self) {
        unsafe { ffi::psd_sponge_free(self.h) }
    }
}

impl<P: FpConfig<4>> Clone for B200PoseidonSponge<P> {
    // #[derive(Clone)] on PoseidonSponge, poseidon/mod.rs:44
    fn clone(&self) -> Self {
        let mut h = std::ptr::null_mut();
        check(unsafe { ffi::psd_sponge_clone(self.h, &mut h) });
        Self { parameters: self.parameters.clone(), ctx: self.ctx.clone(), h }
    }
}

impl<P: FpConfig<4>> B200PoseidonSponge<P> {
    /// Several sponges over the same parameters should share one context.
    pub fn with_context(parameters: &PoseidonConfig<Fp<P, 4>>, ctx: Arc<Ctx>) -> Self {
        let mut h = std::ptr::null_mut();
        check(unsafe { ffi::psd_sponge_new(ctx.raw(), &mut h) });
        Self { parameters: parameters.clone(), ctx, h }
    }
}

impl<P: FpConfig<4>> CryptographicSponge for B200PoseidonSponge<P> {
    type Config = PoseidonConfig<Fp<P, 4>>;

    /// poseidon/mod.rs:219-230
    fn new(parameters: &Self::Config) -> Self {
        Self::with_context(parameters, Ctx::cached(parameters, 0))
    }

    /// poseidon/mod.rs:232-254; the marshalling of `input` is the reference's own (absorb.rs)
    fn absorb(&mut self, input: &impl Absorb) {
        let elems: Vec<Fp<P, 4>> = input.to_sponge_field_elements_as_vec::<Fp<P, 4>>();
        if elems.is_empty() {
            return;
        }
        check(unsafe { ffi::psd_sponge_absorb(self.h, elems.as_ptr() as *const u64, elems.len(), ffi::REPR_MONTGOMERY) });
    }

    /// poseidon/mod.rs:256-270
    fn squeeze_bytes(&mut self, num_bytes: usize) -> Vec<u8> {
        let mut out = vec![0u8; num_bytes];
        check(unsafe { ffi::psd_sponge_squeeze_bytes(self.h, num_bytes, out.as_mut_ptr()) });
        out
    }

    /// poseidon/mod.rs:272-286
    fn squeeze_bits(&mut self, num_bits: usize) -> Vec<bool> {
        let mut raw = vec![0u8; num_bits];
        check(unsafe { ffi::psd_sponge_squeeze_bits(self.h, num_bits, raw.as_mut_ptr()) });
        raw.into_iter().map(|b| b != 0).collect()
    }

    /// poseidon/mod.rs:288-304 -- same body as the reference, on top of the native / bit squeezes above
    fn squeeze_field_elements_with_sizes<F2: PrimeField>(&mut self, sizes: &[FieldElementSize]) -> Vec<F2> {
        if <Fp<P, 4> as PrimeField>::MODULUS.to_bytes_le() == F2::MODULUS.to_bytes_le() {
            self.squeeze_native_field_elements_with_sizes(sizes)
                .iter()
                .map(|x| F2::from_le_bytes_mod_order(&x.into_bigint().to_bytes_le()))
                .collect()
        } else {
            // lib.rs:61-100 (the crate-private default implementation), restated
            let bits_per = (F2::MODULUS_BIT_SIZE - 1) as usize; // FieldElementSize::num_bits always returns this (lib.rs:45-52)
            let bits = self.squeeze_bits(bits_per * sizes.len());
            bits.chunks(bits_per)
                .map(|w| {
                    let bytes: Vec<u8> = w.chunks(8).map(|c| c.iter().enumerate().fold(0u8, |a, (i, &b)| a | ((b as u8) << i))).collect();
                    F2::from_le_bytes_mod_order(&bytes)
                })
                .collect()
        }
    }
    // squeeze_field_elements and fork keep the trait's provided bodies (lib.rs:143-157).
}

impl<P: FpConfig<4>> FieldBasedCryptographicSponge<Fp<P, 4>> for B200PoseidonSponge<P> {
    /// poseidon/mod.rs:320-342
    fn squeeze_native_field_elements(&mut self, num_elements: usize) -> Vec<Fp<P, 4>> {
        let mut out = vec![<Fp<P, 4> as ark_ff::Zero>::zero(); num_elements];
        check(unsafe { ffi::psd_sponge_squeeze_native(self.h, num_elements, out.as_mut_ptr() as *mut u64, ffi::REPR_MONTGOMERY) });
        out
    }
}

impl<P: FpConfig<4>> SpongeExt for B200PoseidonSponge<P> {
    type State = PoseidonSpongeState<Fp<P, 4>>;

    /// poseidon/mod.rs:354-359.  PoseidonSpongeState's fields are private in the reference; this needs the
    /// two accessors (or `pub(crate)` -> `pub`) a maintainer would add alongside this crate.
    fn from_state(state: Self::State, params: &Self::Config) -> Self {
        let s = Self::new(params);
        let (elems, mode) = state.into_parts();
        let (tag, idx) = match mode {
            DuplexSpongeMode::Absorbing { next_absorb_index } => (ffi::MODE_ABSORBING, next_absorb_index),
            DuplexSpongeMode::Squeezing { next_squeeze_index } => (ffi::MODE_SQUEEZING, next_squeeze_index),
        };
        check(unsafe { ffi::psd_sponge_set_state(s.h, elems.as_ptr() as *const u64, ffi::REPR_MONTGOMERY, tag, idx as u32) });
        s
    }

    /// poseidon/mod.rs:361-366
    fn into_state(self) -> Self::State {
        let t = self.parameters.rate + self.parameters.capacity;
        let mut elems = vec![<Fp<P, 4> as ark_ff::Zero>::zero(); t];
        let (mut tag, mut idx) = (0 as c_int, 0u32);
        check(unsafe { ffi::psd_sponge_get_state(self.h, elems.as_mut_ptr() as *mut u64, ffi::REPR_MONTGOMERY, &mut tag, &mut idx) });
        let mode = if tag == ffi::MODE_ABSORBING {
            DuplexSpongeMode::Absorbing { next_absorb_index: idx as usize }
        } else {
            DuplexSpongeMode::Squeezing { next_squeeze_index: idx as usize }
        };
        PoseidonSpongeState::from_parts(elems, mode)
    }
}
