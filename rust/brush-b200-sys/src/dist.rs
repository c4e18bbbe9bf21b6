//! Safe wrappers of the view-parallel (multi-GPU) entry points.  The reference is single-device (train.rs:216-219
//! asserts one view), so none of these has a counterpart there; DESIGN.md section 7 describes the four exchange modes.
//! Peer-mapped buffers (NVLink symmetric memory, NVSwitch multicast addresses) are the caller's to map; the library
//! only dereferences them.  Source only; never compiled here.
use std::os::raw::{c_int, c_void};

use crate::ffi;
use crate::train::AdamHyper;
use crate::{Ctx, Result};

pub use ffi::{BrushDpSegment, BrushDpStep, BrushViewExchange};

/// The slice [begin, end) of a `count`-element segment that `rank` of `world` owns in `dp_adam_step`.
pub fn dp_owned_range(count: u64, world: i32, rank: i32) -> (u64, u64) {
    let (mut b, mut e) = (0u64, 0u64);
    unsafe { ffi::brush_dp_owned_range(count, world as c_int, rank as c_int, &mut b, &mut e) };
    (b, e)
}

impl Ctx {
    /// `views` exchange, publishing half: rasterize_backwards of this rank's view straight into its peer-visible row
    /// block (48 B per splat), visibility flags, the view's uniforms and loss term, then the epoch into every rank's flags.
    ///
    /// # Safety
    /// `x` holds peer-mapped row and flag blocks of every rank; `aux` / `out_img` from this view's forward.
    pub unsafe fn render_backward_rows(
        &self,
        n: u32,
        k: u32,
        out_img: *const f32,
        v_out_img: *const f32,
        aux: &ffi::BrushAux,
        loss_term: *const f32,
        x: &BrushViewExchange,
        stream: *mut c_void,
    ) -> Result<()> {
        self.check(ffi::brush_render_backward_rows(self.as_ptr(), n, k, out_img, v_out_img, aux, loss_term, x, stream))
    }

    /// `views` exchange, consuming half: per-splat backward of EVERY rank's view from the published rows + Adam on this
    /// rank's replica + refine statistics + mean loss, one kernel; replicas stay bit-identical (fixed view order).
    ///
    /// # Safety
    /// As [`Ctx::render_backward_rows`]; `state` is this rank's full replica with its moments.
    pub unsafe fn views_adam_step(
        &self,
        n: u32,
        k: u32,
        state: &ffi::BrushTrainState,
        quats: *const f32,
        x: &BrushViewExchange,
        grad_norm_accum: *mut f32,
        visible_accum: *mut f32,
        loss_out: *mut f32,
        h: AdamHyper,
        stream: *mut c_void,
    ) -> Result<()> {
        self.check(ffi::brush_views_adam_step(
            self.as_ptr(),
            n,
            k,
            state,
            quats,
            x,
            grad_norm_accum,
            visible_accum,
            loss_out,
            h.beta1,
            h.beta2,
            h.eps,
            h.step,
            h.grad_scale,
            stream,
        ))
    }

    /// `hybrid` exchange, publishing half: dense gradients of the eleven small floats per splat + this view's colour
    /// rows (v_r, v_g, v_b, visible: 16 B per splat) into peer-visible memory.
    ///
    /// # Safety
    /// As [`Ctx::render_backward_rows`]; the five gradient arrays hold 3n, 2n, 3n, 4n and n floats.
    pub unsafe fn render_backward_hybrid(
        &self,
        n: u32,
        k: u32,
        p: crate::SplatPtrs,
        out_img: *const f32,
        v_out_img: *const f32,
        aux: &ffi::BrushAux,
        v_means: *mut f32,
        v_xy: *mut f32,
        v_log_scales: *mut f32,
        v_quats: *mut f32,
        v_raw_opacity: *mut f32,
        loss_term: *const f32,
        x: &BrushViewExchange,
        stream: *mut c_void,
    ) -> Result<()> {
        self.check(ffi::brush_render_backward_hybrid(
            self.as_ptr(),
            n,
            k,
            p.means,
            p.log_scales,
            p.quats,
            p.raw_opacity,
            out_img,
            v_out_img,
            aux,
            v_means,
            v_xy,
            v_log_scales,
            v_quats,
            v_raw_opacity,
            loss_term,
            x,
            stream,
        ))
    }

    /// `hybrid` exchange, SH half: rebuilds sum over views of basis(viewdir) (x) v_rgb from every rank's colour rows and
    /// applies Adam to this rank's full SH replica.  `means` = the means the step's forwards rendered with.
    ///
    /// # Safety
    /// sh_coeffs and its two moment arrays hold 3kn floats; `x` as in [`Ctx::render_backward_hybrid`].
    pub unsafe fn sh_views_adam_step(
        &self,
        n: u32,
        k: u32,
        means: *const f32,
        sh_coeffs: *mut f32,
        m_sh_coeffs: *mut f32,
        v_sh_coeffs: *mut f32,
        lr_coeffs_dc: f32,
        sh_rest_lr_scale: f32,
        x: &BrushViewExchange,
        loss_out: *mut f32,
        h: AdamHyper,
        stream: *mut c_void,
    ) -> Result<()> {
        self.check(ffi::brush_sh_views_adam_step(
            self.as_ptr(),
            n,
            k,
            means,
            sh_coeffs,
            m_sh_coeffs,
            v_sh_coeffs,
            lr_coeffs_dc,
            sh_rest_lr_scale,
            x,
            loss_out,
            h.beta1,
            h.beta2,
            h.eps,
            h.step,
            h.grad_scale,
            stream,
        ))
    }

    /// `peer` exchange: sum of the ranks' dense gradients on the owned slice (peer loads or ONE multimem.ld_reduce) ->
    /// Adam (moments sharded) -> new parameters into every replica (peer stores or multimem.st), one kernel.
    ///
    /// # Safety
    /// `step` holds peer-mapped gradient / parameter / flag blocks of every rank (and their multicast addresses, or null).
    pub unsafe fn dp_adam_step(
        &self,
        step: &BrushDpStep,
        segments: &[BrushDpSegment],
        h: AdamHyper,
        stream: *mut c_void,
    ) -> Result<()> {
        self.check(ffi::brush_dp_adam_step(
            self.as_ptr(),
            step,
            segments.as_ptr(),
            segments.len() as c_int,
            h.beta1,
            h.beta2,
            h.eps,
            h.step,
            h.grad_scale,
            stream,
        ))
    }

    /// The NEXT forward on this context reads `sh_coeffs` only behind `event` (a `cudaEvent_t`; one-shot) and evaluates
    /// colours in their own kernel just before the rasterizer, so projection / sorts / binning run under an SH exchange
    /// still in flight on another stream.  Null clears it.
    pub fn set_sh_ready_event(&self, event: *mut c_void) -> Result<()> {
        self.check(unsafe { ffi::brush_ctx_set_sh_ready_event(self.as_ptr(), event) })
    }

    /// Always run the deferred-colour forward (A/B measurements; same bits as the fused forward).
    pub fn set_split_colors(&self, enabled: bool) -> Result<()> {
        self.check(unsafe { ffi::brush_ctx_set_split_colors(self.as_ptr(), enabled as c_int) })
    }
}
