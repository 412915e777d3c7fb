//! Safe, slice-level wrapper over `b200tfhe-sys`.
//!
//! The shapes follow the reference functions this library replaces:
//! `keyswitch_lwe_ciphertext` (core_crypto/algorithms/lwe_keyswitch.rs:96-170),
//! `programmable_bootstrap_lwe_ciphertext_mem_optimized` (lwe_programmable_bootstrapping.rs:1067-1113) and shortint's
//! `keyswitch_programmable_bootstrap_assign` (shortint/server_key/mod.rs:783-857), over a LIST of ciphertexts: the containers are
//! flat `&[u64]` slices with the reference's memory layout (mask then body), `batch` ciphertexts back to back.
//!
//! Not compiled in the build image of this repository (it has no Rust toolchain); the declarations it relies on are generated
//! from the C header and checked by the test-suite, and the same call sequence is exercised from C by
//! tests/c_api/test_shortint_ks_pbs.c and from Python by the ctypes mirror.
use b200tfhe_sys as sys;
use std::ffi::CStr;
use std::ptr;

#[derive(Debug)]
pub struct Error(pub String);
pub type Result<T> = std::result::Result<T, Error>;

pub use sys::b200tfhe_params as Params;

/// One CUDA device + stream + workspace pool; one per host thread, like the reference's thread-local engines.
pub struct Context {
    raw: *mut sys::b200tfhe_context,
}
unsafe impl Send for Context {}

impl Context {
    pub fn new(device: i32) -> Result<Self> {
        let mut raw = ptr::null_mut();
        let rc = unsafe { sys::b200tfhe_context_create(device, &mut raw) };
        if rc != 0 || raw.is_null() {
            return Err(Error(format!("b200tfhe_context_create({device}) failed")));
        }
        Ok(Self { raw })
    }
    fn check(&self, rc: i32) -> Result<()> {
        if rc == 0 {
            return Ok(());
        }
        let msg = unsafe { CStr::from_ptr(sys::b200tfhe_last_error(self.raw)) };
        Err(Error(msg.to_string_lossy().into_owned()))
    }
    pub fn set_option(&self, key: &CStr, value: i64) -> Result<()> {
        self.check(unsafe { sys::b200tfhe_context_set_option(self.raw, key.as_ptr(), value) })
    }
}
impl Drop for Context {
    fn drop(&mut self) {
        unsafe { sys::b200tfhe_context_destroy(self.raw) };
    }
}

/// Device-resident Fourier bootstrap key (`FourierLweBootstrapKey`); borrows the context it was created on.
pub struct FourierBsk<'c> {
    ctx: &'c Context,
    raw: *mut sys::b200tfhe_bsk,
    pub params: Params,
}
/// Device-resident keyswitch key (`LweKeyswitchKey`).
pub struct Ksk<'c> {
    ctx: &'c Context,
    raw: *mut sys::b200tfhe_ksk,
    pub params: Params,
}

impl<'c> FourierBsk<'c> {
    /// `par_convert_standard_lwe_bootstrap_key_to_fourier`: `std_bsk` is the standard-domain key `[n][l][k+1][k+1][N]`.
    pub fn from_standard(ctx: &'c Context, params: Params, std_bsk: &[u64]) -> Result<Self> {
        let (n, k, nn, l) = (params.lwe_dimension as usize, params.glwe_dimension as usize, params.polynomial_size as usize, params.pbs_level as usize);
        if std_bsk.len() != n * l * (k + 1) * (k + 1) * nn {
            return Err(Error("bootstrap key: wrong number of scalars".into()));
        }
        let mut raw = ptr::null_mut();
        ctx.check(unsafe { sys::b200tfhe_bsk_upload_u64(ctx.raw, &params, std_bsk.as_ptr(), &mut raw) })?;
        Ok(Self { ctx, raw, params })
    }
}
impl Drop for FourierBsk<'_> {
    fn drop(&mut self) {
        unsafe { sys::b200tfhe_bsk_destroy(self.raw) };
    }
}
impl<'c> Ksk<'c> {
    /// `ksk`: `[kN][l_ks][n+1]`, levels stored l_ks..1 (lwe_keyswitch_key_generation.rs:109-120).
    pub fn upload(ctx: &'c Context, params: Params, ksk: &[u64]) -> Result<Self> {
        let (n, k, nn, l) = (params.lwe_dimension as usize, params.glwe_dimension as usize, params.polynomial_size as usize, params.ks_level as usize);
        if ksk.len() != k * nn * l * (n + 1) {
            return Err(Error("keyswitch key: wrong number of scalars".into()));
        }
        let mut raw = ptr::null_mut();
        ctx.check(unsafe { sys::b200tfhe_ksk_upload_u64(ctx.raw, &params, ksk.as_ptr(), &mut raw) })?;
        Ok(Self { ctx, raw, params })
    }
    /// `keyswitch_lwe_ciphertext` over a list: `input` holds `batch` ciphertexts of kN+1 scalars, `output` n+1 each.
    pub fn keyswitch_batch(&self, input: &[u64], output: &mut [u64]) -> Result<()> {
        let big = (self.params.glwe_dimension * self.params.polynomial_size + 1) as usize;
        let small = (self.params.lwe_dimension + 1) as usize;
        if input.len() % big != 0 || output.len() != input.len() / big * small {
            return Err(Error("keyswitch_batch: slice lengths do not match the key dimensions".into()));
        }
        let batch = (input.len() / big) as u32;
        self.ctx.check(unsafe { sys::b200tfhe_keyswitch_batch_u64(self.ctx.raw, self.raw, input.as_ptr(), output.as_mut_ptr(), batch) })
    }
}
impl Drop for Ksk<'_> {
    fn drop(&mut self) {
        unsafe { sys::b200tfhe_ksk_destroy(self.raw) };
    }
}

/// shortint `keyswitch_programmable_bootstrap_assign` core over a list: keyswitch, then programmable bootstrap with accumulator
/// `luts[lut_idx[i]]` (every ciphertext uses `luts[0]` when `lut_idx` is `None`).  `input` / `output`: `batch x (kN+1)`;
/// `luts`: `n_luts x (k+1)N` (GLWE accumulators as built by shortint's `fill_accumulator`).
pub fn keyswitch_programmable_bootstrap_batch(ksk: &Ksk, bsk: &FourierBsk, input: &[u64], output: &mut [u64], luts: &[u64],
                                              lut_idx: Option<&[u32]>) -> Result<()> {
    let p = &bsk.params;
    let big = (p.glwe_dimension * p.polynomial_size + 1) as usize;
    let glwe = ((p.glwe_dimension + 1) * p.polynomial_size) as usize;
    if input.len() % big != 0 || output.len() != input.len() || luts.is_empty() || luts.len() % glwe != 0 {
        return Err(Error("keyswitch_programmable_bootstrap_batch: slice lengths do not match the key dimensions".into()));
    }
    let batch = input.len() / big;
    let n_luts = (luts.len() / glwe) as u32;
    if let Some(idx) = lut_idx {
        if idx.len() != batch || idx.iter().any(|&i| i >= n_luts) {
            return Err(Error("keyswitch_programmable_bootstrap_batch: lut_idx must hold one valid table index per ciphertext".into()));
        }
    }
    let idx_ptr = lut_idx.map_or(ptr::null(), |s| s.as_ptr());
    ksk.ctx.check(unsafe {
        sys::b200tfhe_ks_pbs_batch_u64(ksk.ctx.raw, ksk.raw, bsk.raw, input.as_ptr(), output.as_mut_ptr(), luts.as_ptr(), n_luts, idx_ptr, batch as u32)
    })
}
