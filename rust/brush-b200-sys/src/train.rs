//! Safe wrappers of the train-step entry points: what `SplatTrainer::step` (crates/brush-train/src/train.rs:211-393)
//! and `refine_splats` (train.rs:395-553) do around the render call -- image loss, rotation normalisation, the fused
//! backward + Adam, refine statistics, densify / prune -- plus the remaining primitives and diagnostics.
//! Source only; never compiled here.  Every pointer is a device pointer unless its name ends in `_host`.
use std::ffi::CStr;
use std::os::raw::{c_char, c_int, c_void};

use crate::ffi;
use crate::{Ctx, Error, Result};

pub use ffi::{BrushAdamGroup, BrushRefineConfig, BrushRefineCounts, BrushSplatParams, BrushTrainState};

/// Adam hyper-parameters as brush-train sets them (train.rs:184: beta 0.9 / 0.999, eps 1e-15) and the step number of
/// the bias correction, which restarts at 1 after every refine (train.rs:568 re-creates the optimizer).
#[derive(Clone, Copy, Debug)]
pub struct AdamHyper {
    pub beta1: f32,
    pub beta2: f32,
    pub eps: f32,
    pub step: u32,
    /// 1 / (views in the global batch): the mean over the batch of train.rs:229-268.
    pub grad_scale: f32,
}

impl Default for AdamHyper {
    fn default() -> Self {
        Self { beta1: 0.9, beta2: 0.999, eps: 1e-15, step: 1, grad_scale: 1.0 }
    }
}

/// `brush_version()`: "brush_b200 <version> sm_100a".
pub fn version() -> String {
    unsafe { CStr::from_ptr(ffi::brush_version()) }.to_string_lossy().into_owned()
}

/// Kernels launched by the library in this process, all contexts (the bench's `gpu_launches`).
pub fn launch_count() -> u64 {
    unsafe { ffi::brush_launch_count() }
}

/// The allocation-free scan (`brush_prefix_sum_u32_noalloc`; no counterpart in the reference): no context, scratch in
/// a static device array inside the library, at most `ffi::BRUSH_NOALLOC_SCAN_MAX_N` elements, current device.  Calls
/// on one device share the scratch: keep them on one stream or order them.
///
/// # Safety
/// `input` and `out` hold `n` elements on the current device.
pub unsafe fn prefix_sum_u32_noalloc(input: *const u32, n: u32, out: *mut u32, stream: *mut c_void) -> Result<()> {
    let mut err = [0 as c_char; 256];
    let rc = ffi::brush_prefix_sum_u32_noalloc(input, n, out, stream, err.as_mut_ptr(), err.len() as u32);
    let text = || CStr::from_ptr(err.as_ptr()).to_string_lossy().into_owned();
    match rc {
        ffi::BRUSH_OK => Ok(()),
        ffi::BRUSH_E_INVALID => Err(Error::Invalid(text())),
        _ => Err(Error::Cuda(text())),
    }
}

impl Ctx {
    /// 64-bit-key variant of [`Ctx::radix_argsort_u32`] (the fused tile | depth key of render.rs:227-237).
    ///
    /// # Safety
    /// All four arrays hold `max_n` elements on this context's device.
    pub unsafe fn radix_argsort_u64(
        &self,
        keys: *const u64,
        values: *const u32,
        n_dev: *const i32,
        max_n: u32,
        bits: u32,
        keys_out: *mut u64,
        values_out: *mut u32,
        stream: *mut c_void,
    ) -> Result<()> {
        self.check(ffi::brush_radix_argsort_u64(
            self.as_ptr(), keys, values, n_dev, max_n, bits, keys_out, values_out, stream,
        ))
    }

    /// Host-staged forward for callers that hold HOST buffers: copies the five parameter arrays to the device,
    /// renders, copies the image and the two counters back, synchronises.  Returns (num_visible, num_intersections).
    ///
    /// # Safety
    /// Host slices of the documented lengths: means / log_scales 3n, quats 4n, sh_coeffs 3kn, raw_opacity n,
    /// out_img 4 * width * height.
    pub unsafe fn render_forward_host(
        &self,
        uniforms: &ffi::BrushUniforms,
        n: u32,
        k: u32,
        means: &[f32],
        log_scales: &[f32],
        quats: &[f32],
        sh_coeffs: &[f32],
        raw_opacity: &[f32],
        isect_capacity: u64,
        out_img: &mut [f32],
    ) -> Result<(u32, u32)> {
        let (mut visible, mut intersections) = (0u32, 0u32);
        self.check(ffi::brush_render_forward_host(
            self.as_ptr(),
            uniforms,
            n,
            k,
            means.as_ptr(),
            log_scales.as_ptr(),
            quats.as_ptr(),
            sh_coeffs.as_ptr(),
            raw_opacity.as_ptr(),
            isect_capacity,
            out_img.as_mut_ptr(),
            &mut visible,
            &mut intersections,
        ))
        .map(|_| (visible, intersections))
    }

    /// u8 -> f32 / 255 of an uploaded view image (the reference's `image_to_tensor`), on the device.
    ///
    /// # Safety
    /// `src` holds `n` bytes and `dst` `n` floats on this context's device.
    pub unsafe fn image_u8_to_f32(&self, src: *const u8, n: u64, dst: *mut f32, stream: *mut c_void) -> Result<()> {
        self.check(ffi::brush_image_u8_to_f32(self.as_ptr(), src, n, dst, stream))
    }

    /// `(pred - gt).abs().mean()` and its gradient (train.rs:249-255).
    ///
    /// # Safety
    /// pred f32[H,W,4], gt f32[H,W,gt_channels], v_out_img f32[H,W,4], loss_out f32[1], all on the device.
    pub unsafe fn l1_loss_backward(
        &self,
        pred: *const f32,
        gt: *const f32,
        height: u32,
        width: u32,
        gt_channels: u32,
        v_out_img: *mut f32,
        loss_out: *mut f32,
        stream: *mut c_void,
    ) -> Result<()> {
        self.check(ffi::brush_l1_loss_backward(
            self.as_ptr(), pred, gt, height, width, gt_channels, v_out_img, loss_out, stream,
        ))
    }

    /// `(1 - w) * L1 - w * ssim(pred, gt).mean()` and d loss / d pred in two launches (train.rs:249-268, ssim.rs).
    /// `ssim_window` must be 11 (the reference default, train.rs:59-60).
    ///
    /// # Safety
    /// As [`Ctx::l1_loss_backward`].
    pub unsafe fn image_loss_backward(
        &self,
        pred: *const f32,
        gt: *const f32,
        height: u32,
        width: u32,
        gt_channels: u32,
        ssim_weight: f32,
        ssim_window: u32,
        v_out_img: *mut f32,
        loss_out: *mut f32,
        stream: *mut c_void,
    ) -> Result<()> {
        self.check(ffi::brush_image_loss_backward(
            self.as_ptr(), pred, gt, height, width, gt_channels, ssim_weight, ssim_window, v_out_img, loss_out, stream,
        ))
    }

    /// Rotation normalisation of `Splats::render` (gaussian_splats.rs:174-175).
    ///
    /// # Safety
    /// `quats` and `out` hold 4n floats on this context's device.
    pub unsafe fn quat_normalize(&self, quats: *const f32, n: u32, out: *mut f32, stream: *mut c_void) -> Result<()> {
        self.check(ffi::brush_quat_normalize(self.as_ptr(), quats, n, out, stream))
    }

    /// VJP of the normalisation, in place on the gradient of the normalised rotations (Burn autodiff in the reference).
    ///
    /// # Safety
    /// `quats` (the raw rotations) and `v_inout` hold 4n floats on this context's device.
    pub unsafe fn quat_normalize_vjp(
        &self,
        quats: *const f32,
        n: u32,
        v_inout: *mut f32,
        stream: *mut c_void,
    ) -> Result<()> {
        self.check(ffi::brush_quat_normalize_vjp(self.as_ptr(), quats, n, v_inout, stream))
    }

    /// Refine statistics of one view (train.rs:284-316): |v_xy * (w/2, h/2)| and the visibility count per splat.
    ///
    /// # Safety
    /// v_xy f32[n,2], grad_norm / visible f32[n]; `aux` from the view's forward.
    pub unsafe fn xy_grad_stats(
        &self,
        v_xy: *const f32,
        aux: &ffi::BrushAux,
        n: u32,
        width: u32,
        height: u32,
        accumulate: bool,
        grad_norm: *mut f32,
        visible: *mut f32,
        stream: *mut c_void,
    ) -> Result<()> {
        self.check(ffi::brush_xy_grad_stats(
            self.as_ptr(), v_xy, aux, n, width, height, accumulate as c_int, grad_norm, visible, stream,
        ))
    }

    /// The five `optim.step` calls of train.rs:318-359 as one launch (Burn Adam, eps 1e-15).
    ///
    /// # Safety
    /// Every group's four arrays hold `count` floats on this context's device.
    pub unsafe fn adam_step(&self, groups: &[BrushAdamGroup], h: AdamHyper, stream: *mut c_void) -> Result<()> {
        self.check(ffi::brush_adam_step(
            self.as_ptr(), groups.as_ptr(), groups.len() as c_int, h.beta1, h.beta2, h.eps, h.step, h.grad_scale, stream,
        ))
    }

    /// `RenderBackwards::backward` + the normalisation VJP + the five optimizer steps + the refine statistics for the
    /// reference's own shape (one view, one device): rasterize_backwards and ONE per-splat kernel; the dense gradients
    /// never exist.  `grad_norm_accum` / `visible_accum` may be null (train.rs:286: statistics start after warm-up).
    ///
    /// # Safety
    /// `state` holds the parameters and moments of all five groups; `quats` the normalised rotations the forward used.
    pub unsafe fn render_backward_adam(
        &self,
        n: u32,
        k: u32,
        state: &BrushTrainState,
        quats: *const f32,
        out_img: *const f32,
        v_out_img: *const f32,
        aux: &ffi::BrushAux,
        v_xy: *mut f32,
        grad_norm_accum: *mut f32,
        visible_accum: *mut f32,
        h: AdamHyper,
        stream: *mut c_void,
    ) -> Result<()> {
        self.check(ffi::brush_render_backward_adam(
            self.as_ptr(),
            n,
            k,
            state,
            quats,
            out_img,
            v_out_img,
            aux,
            v_xy,
            grad_norm_accum,
            visible_accum,
            h.beta1,
            h.beta2,
            h.eps,
            h.step,
            h.grad_scale,
            stream,
        ))
    }

    /// First half of `refine_splats` (train.rs:395-553): classifies every splat (clone / split / prune) and returns the
    /// `RefineStats` counts and the size of the new set.  Synchronises, as the reference awaits its `argwhere`s.
    ///
    /// # Safety
    /// The accumulators hold `n` floats; `pre_step` / `post_step` describe `n` splats.
    pub unsafe fn refine_plan(
        &self,
        n: u32,
        grad_2d_accum: *const f32,
        xy_grad_counts: *const f32,
        pre_step: &BrushSplatParams,
        post_step: &BrushSplatParams,
        config: &BrushRefineConfig,
        stream: *mut c_void,
    ) -> Result<BrushRefineCounts> {
        let mut counts: BrushRefineCounts = std::mem::zeroed();
        self.check(ffi::brush_refine_plan(
            self.as_ptr(), n, grad_2d_accum, xy_grad_counts, pre_step, post_step, config, &mut counts, stream,
        ))
        .map(|_| counts)
    }

    /// Second half: writes the new set (kept, cloned, split; `prune_points` train.rs:560-596 and `concat_splats`
    /// :598-621) into caller-allocated `out` in the reference's order.  `samples` = the N(0, 0.5) draws of the splits.
    ///
    /// # Safety
    /// Must follow a `refine_plan` of the same `n`; `out` holds the count that plan returned.
    pub unsafe fn refine_apply(
        &self,
        n: u32,
        k: u32,
        pre_step: &BrushSplatParams,
        post_step: &BrushSplatParams,
        samples: *const f32,
        config: &BrushRefineConfig,
        out: &BrushSplatParams,
        stream: *mut c_void,
    ) -> Result<()> {
        self.check(ffi::brush_refine_apply(self.as_ptr(), n, k, pre_step, post_step, samples, config, out, stream))
    }

    /// Measured roofline denominators of this device: [FFMA with constant operands, FFMA with three register operands,
    /// packed FFMA2] in TFLOP/s and MUFU.EX2 in 1e12 / s.
    pub fn measure_peaks(&self, stream: *mut c_void) -> Result<[f32; 4]> {
        let mut out = [0f32; 4];
        self.check(unsafe { ffi::brush_measure_peaks(self.as_ptr(), out.as_mut_ptr(), stream) }).map(|_| out)
    }

    /// Diagnostic: the rasterizers count the (pixel, splat) pairs they evaluate after culling.  Never on in a timed region.
    pub fn set_pair_counting(&self, enabled: bool) -> Result<()> {
        self.check(unsafe { ffi::brush_ctx_set_pair_counting(self.as_ptr(), enabled as c_int) })
    }

    /// Synchronises; returns [forward, backward] pair counts since the last read and clears them.
    pub fn read_pair_counts(&self, stream: *mut c_void) -> Result<[u64; 2]> {
        let mut out = [0u64; 2];
        self.check(unsafe { ffi::brush_ctx_read_pair_counts(self.as_ptr(), out.as_mut_ptr(), stream) }).map(|_| out)
    }
}
