//! `brush-b200-sys`: the raw FFI surface of `libbrush_b200.so` ([`ffi`], generated from `include/brush_b200.h` by
//! `rust/gen_ffi.py`) plus the thin safe layer every caller needs: an owned context, error mapping, and typed
//! argument packs for the two operator entry points that replace
//! `brush_render::render::render_forward` (crates/brush-render/src/render.rs:55-323) and
//! `RenderBackwards::backward` (render.rs:465-626).
//!
//! Source only -- never compiled here (no Rust toolchain in the build image).  The Burn-facing glue that calls this
//! crate is `rust/brush-render-b200/render_b200.rs`.  [`train`] wraps the train-step entry points (image loss, fused
//! backward + Adam, refine), [`dist`] the view-parallel ones; between the three modules every entry point of the header has
//! a safe wrapper (tests/test_abi_host.py checks that).
pub mod dist;
pub mod ffi;
pub mod train;

use std::ffi::CStr;
use std::os::raw::{c_int, c_void};
use std::ptr::NonNull;

pub use ffi::{BrushAux, BrushUniforms};

/// Error codes of the C ABI (`BRUSH_E_*`).  The reference panics on bad dimensions (`DimCheck`, render.rs:74-79;
/// "Invalid nr. of sh bases", render.rs:51) and truncates silently on intersection overflow (render.rs:204-206); the
/// library reports both.
#[derive(Debug, Clone, PartialEq, Eq)]
pub enum Error {
    Invalid(String),
    Capacity(String),
    Cuda(String),
}

impl std::fmt::Display for Error {
    fn fmt(&self, f: &mut std::fmt::Formatter<'_>) -> std::fmt::Result {
        match self {
            Error::Invalid(m) => write!(f, "brush_b200: invalid argument: {m}"),
            Error::Capacity(m) => write!(f, "brush_b200: intersection capacity exceeded: {m}"),
            Error::Cuda(m) => write!(f, "brush_b200: CUDA error: {m}"),
        }
    }
}
impl std::error::Error for Error {}

pub type Result<T> = std::result::Result<T, Error>;

/// One `brush_ctx` per (device, stream): owns scratch only (sort ping-pong, look-back status words, the per-splat
/// gradient rows).  Not `Sync`: the C context is not thread-safe.
pub struct Ctx {
    raw: NonNull<ffi::brush_ctx>,
    device: i32,
}

unsafe impl Send for Ctx {}

impl Ctx {
    pub fn new(device: i32) -> Result<Self> {
        let mut raw: *mut ffi::brush_ctx = std::ptr::null_mut();
        let rc = unsafe { ffi::brush_ctx_create(device as c_int, &mut raw) };
        match NonNull::new(raw) {
            Some(raw) if rc == ffi::BRUSH_OK => Ok(Self { raw, device }),
            _ => Err(Error::Cuda(format!("brush_ctx_create({device}) returned {rc}"))),
        }
    }

    pub fn device(&self) -> i32 {
        self.device
    }

    pub fn as_ptr(&self) -> *mut ffi::brush_ctx {
        self.raw.as_ptr()
    }

    fn last_error(&self) -> String {
        unsafe { CStr::from_ptr(ffi::brush_last_error(self.raw.as_ptr())) }.to_string_lossy().into_owned()
    }

    /// Maps a return code to `Result`.
    pub fn check(&self, rc: c_int) -> Result<()> {
        match rc {
            ffi::BRUSH_OK => Ok(()),
            ffi::BRUSH_E_INVALID => Err(Error::Invalid(self.last_error())),
            ffi::BRUSH_E_CAPACITY => Err(Error::Capacity(self.last_error())),
            _ => Err(Error::Cuda(self.last_error())),
        }
    }

    /// Synchronises `stream` and surfaces the sticky capacity-overflow bit of every forward since the last call.
    pub fn sync_check(&self, stream: *mut c_void) -> Result<()> {
        self.check(unsafe { ffi::brush_ctx_check(self.raw.as_ptr(), stream) })
    }

    /// Tile-sort strategy: `false` = depth sort, then stable sort by tile id (the reference's structure,
    /// render.rs:151-156,225-237); `true` = one fused 64-bit (tile | depth) key sort.  Same output.
    pub fn set_fused_sort(&self, fused64: bool) -> Result<()> {
        self.check(unsafe { ffi::brush_ctx_set_sort_mode(self.raw.as_ptr(), fused64 as c_int) })
    }

    /// Per-stage CUDA-event timing == the `sync_burn` spans of render.rs (crates/sync-span/src/lib.rs:26-41).
    pub fn set_timing(&self, enabled: bool) -> Result<()> {
        self.check(unsafe { ffi::brush_ctx_set_timing(self.raw.as_ptr(), enabled as c_int) })
    }

    pub fn stage_times_ms(&self, stream: *mut c_void) -> Result<[f32; ffi::BRUSH_STAGE_COUNT]> {
        let mut out = [0f32; ffi::BRUSH_STAGE_COUNT];
        let rc = unsafe {
            ffi::brush_stage_times_ms(self.raw.as_ptr(), out.as_mut_ptr(), ffi::BRUSH_STAGE_COUNT as c_int, stream)
        };
        self.check(rc).map(|_| out)
    }
}

impl Drop for Ctx {
    fn drop(&mut self) {
        unsafe { ffi::brush_ctx_destroy(self.raw.as_ptr()) }
    }
}

/// Device pointers of the five parameter arrays as `render_forward` takes them (render.rs:55-64): means [N,3],
/// log_scales [N,3], quats [N,4] (already normalised, gaussian_splats.rs:174-175), sh_coeffs [N,K,3], raw_opacity [N].
#[derive(Clone, Copy)]
pub struct SplatPtrs {
    pub means: *const f32,
    pub log_scales: *const f32,
    pub quats: *const f32,
    pub sh_coeffs: *const f32,
    pub raw_opacity: *const f32,
}

/// Device pointers of the six dense gradients in the registration order of render.rs:598-624.
#[derive(Clone, Copy)]
pub struct GradPtrs {
    pub v_means: *mut f32,
    pub v_xy: *mut f32,
    pub v_log_scales: *mut f32,
    pub v_quats: *mut f32,
    pub v_sh_coeffs: *mut f32,
    pub v_raw_opacity: *mut f32,
}

impl Ctx {
    /// == `render_forward` (render.rs:55-323).  Enqueues on `stream` and returns.
    ///
    /// # Safety
    /// Every pointer must be a live device allocation of the documented size on this context's device.
    pub unsafe fn render_forward(
        &self,
        uniforms: &BrushUniforms,
        n: u32,
        k: u32,
        p: SplatPtrs,
        raster_u32: bool,
        out_img: *mut c_void,
        aux: &BrushAux,
        stream: *mut c_void,
    ) -> Result<()> {
        self.check(ffi::brush_render_forward(
            self.raw.as_ptr(),
            uniforms,
            n,
            k,
            p.means,
            p.log_scales,
            p.quats,
            p.sh_coeffs,
            p.raw_opacity,
            raster_u32 as c_int,
            out_img,
            aux,
            stream,
        ))
    }

    /// == `RenderBackwards::backward` (render.rs:468-625): the three dispatches (RasterizeBackwards, GatherGrads,
    /// ProjectBackwards) and their nine zero-fills as one call.  The six outputs are fully written (zero for splats
    /// no pixel saw), so the caller allocates them uninitialised.
    ///
    /// # Safety
    /// As [`Ctx::render_forward`]; `aux` and `out_img` must come from the matching forward.
    pub unsafe fn render_backward(
        &self,
        n: u32,
        k: u32,
        p: SplatPtrs,
        out_img: *const f32,
        v_out_img: *const f32,
        aux: &BrushAux,
        g: GradPtrs,
        stream: *mut c_void,
    ) -> Result<()> {
        self.check(ffi::brush_render_backward(
            self.raw.as_ptr(),
            n,
            k,
            p.means,
            p.log_scales,
            p.quats,
            p.raw_opacity,
            out_img,
            v_out_img,
            aux,
            g.v_means,
            g.v_xy,
            g.v_log_scales,
            g.v_quats,
            g.v_sh_coeffs,
            g.v_raw_opacity,
            stream,
        ))
    }

    /// == `brush_sort::radix_argsort` (crates/brush-sort/src/lib.rs:32-37): stable, first `*n_dev` entries, the low
    /// `4 * ceil(bits / 4)` key bits.
    ///
    /// # Safety
    /// All four arrays hold `max_n` elements on this context's device.
    pub unsafe fn radix_argsort_u32(
        &self,
        keys: *const u32,
        values: *const u32,
        n_dev: *const i32,
        max_n: u32,
        bits: u32,
        keys_out: *mut u32,
        values_out: *mut u32,
        stream: *mut c_void,
    ) -> Result<()> {
        self.check(ffi::brush_radix_argsort_u32(
            self.raw.as_ptr(),
            keys,
            values,
            n_dev,
            max_n,
            bits,
            keys_out,
            values_out,
            stream,
        ))
    }

    /// == `brush_prefix_sum::prefix_sum` (crates/brush-prefix-sum/src/lib.rs:17): inclusive u32 scan.
    ///
    /// # Safety
    /// `input` and `out` hold `n` elements on this context's device.
    pub unsafe fn prefix_sum_u32(&self, input: *const u32, n: u32, out: *mut u32, stream: *mut c_void) -> Result<()> {
        self.check(ffi::brush_prefix_sum_u32(self.raw.as_ptr(), input, n, out, stream))
    }
}

/// `min(N * tiles, 128 * 65535)`, the intersection buffer length the reference hard-codes (render.rs:204-206).
pub fn reference_max_intersects(num_points: u32, tile_bounds: [u32; 2]) -> u64 {
    let tiles = tile_bounds[0] as u64 * tile_bounds[1] as u64;
    (num_points as u64 * tiles).min(128 * 65535)
}

#[cfg(test)]
mod tests {
    use super::*;

    #[test]
    fn uniforms_layout_matches_wgsl() {
        // shaders::helpers::RenderUniforms, helpers.wgsl:7-30: 112 bytes, num_visible is word 25 (render.rs:145-149)
        assert_eq!(std::mem::size_of::<BrushUniforms>(), 112);
        let u: BrushUniforms = unsafe { std::mem::zeroed() };
        let base = &u as *const _ as usize;
        assert_eq!(&u.num_visible as *const _ as usize - base, 25 * 4);
    }

    #[test]
    fn max_intersects() {
        assert_eq!(reference_max_intersects(10, [2, 2]), 40);
        assert_eq!(reference_max_intersects(1_000_000, [120, 68]), 128 * 65535);
    }
}
