//! `SplatTrainer::step` / `refine_splats` on the B200-native kernels -- the train-step half of the drop-in (SURVEY.md
//! 8f, row f-1 / f-2), compiled as a module of brush-train behind a `b200` feature next to
//! `brush_render::render_b200` (see ../brush-render-b200/README.md for the feature wiring).
//!
//! Mirrors crates/brush-train/src/train.rs:
//!   :151-198  SplatTrainer::new              -> [`SplatTrainerB200::new`]   (Adam moments as plain tensors)
//!   :211-393  SplatTrainer::step             -> [`SplatTrainerB200::step`]  (render -> image loss -> ONE backward +
//!                                               optimizer kernel; refine statistics come out of the same kernel)
//!   :395-553  refine_splats (+ prune_points :560-596, concat_splats :598-621) -> [`SplatTrainerB200::refine_splats`]
//! What is gone: `loss.backward()` (the autodiff graph of the loss, the render node and the normalisation), the
//! `GradientsParams` plumbing, five `optim.step` calls and the SH lerp hack (train.rs:318-359), the `select_assign` /
//! `powi_scalar` housekeeping (:284-316), and the dozen Burn ops + two `argwhere` readbacks of refine_splats.
//! Source only: never compiled in the build image (no Rust toolchain).
use std::os::raw::c_void;

use brush_b200_sys::train::{AdamHyper, BrushRefineConfig, BrushSplatParams, BrushTrainState};
use brush_b200_sys::{BrushAux, Ctx};
use brush_render::gaussian_splats::{inverse_sigmoid, Splats};
use brush_render::render_b200::{aux_of, ctx_for, dptr, stream_of, CudaBackend};
use brush_render::RenderAux;
use burn::backend::Autodiff;
use burn::lr_scheduler::LrScheduler;
use burn::tensor::{Distribution, Tensor};

use crate::train::{RefineStats, SceneBatch, TrainConfig, TrainStepStats};

type B = Autodiff<CudaBackend>;
type Inner = CudaBackend;

/// First and second Adam moments of one parameter group (Burn keeps them inside `OptimizerAdaptor` records,
/// train.rs:186-193; the fused kernel updates them in place).
struct Moments<const D: usize> {
    m: Tensor<Inner, D>,
    v: Tensor<Inner, D>,
}

impl<const D: usize> Moments<D> {
    fn zeros_like(p: &Tensor<Inner, D>) -> Self {
        Self { m: Tensor::zeros_like(p), v: Tensor::zeros_like(p) }
    }
}

pub struct SplatTrainerB200 {
    pub iter: u32,
    config: TrainConfig,
    sched_mean: <burn::lr_scheduler::exponential::ExponentialLrSchedulerConfig as burn::config::Config>::Scheduler,
    /// Adam's bias-correction time: 1 on the first step after `new` and after every refine (train.rs:568 re-creates
    /// the optimizer, so Burn's per-parameter `time` restarts together with the moments).
    opt_step: u32,
    mom_means: Moments<2>,
    mom_sh: Moments<3>,
    mom_rot: Moments<2>,
    mom_opac: Moments<1>,
    mom_scales: Moments<2>,
    grad_2d_accum: Tensor<Inner, 1>,
    xy_grad_counts: Tensor<Inner, 1>,
    // step scratch, allocated once per image size / splat count
    norm_rot: Tensor<Inner, 2>,
    v_out: Option<Tensor<Inner, 3>>,
    v_xy: Tensor<Inner, 2>,
    loss: Tensor<Inner, 1>,
}

impl SplatTrainerB200 {
    /// == SplatTrainer::new, train.rs:167-198.
    pub fn new(splats: &Splats<B>, config: &TrainConfig) -> Self {
        let device = splats.means.device();
        let n = splats.num_splats();
        let inner = |t: &Tensor<B, 2>| t.clone().inner();
        Self {
            iter: 0,
            config: config.clone(),
            sched_mean: config.lr_mean.init(),
            opt_step: 0,
            mom_means: Moments::zeros_like(&inner(&splats.means.val())),
            mom_sh: Moments::zeros_like(&splats.sh_coeffs.val().inner()),
            mom_rot: Moments::zeros_like(&inner(&splats.rotation.val())),
            mom_opac: Moments::zeros_like(&splats.raw_opacity.val().inner()),
            mom_scales: Moments::zeros_like(&inner(&splats.log_scales.val())),
            grad_2d_accum: Tensor::zeros([n], &device),
            xy_grad_counts: Tensor::zeros([n], &device),
            norm_rot: Tensor::empty([n, 4], &device),
            v_out: None,
            v_xy: Tensor::empty([n, 2], &device),
            loss: Tensor::zeros([1], &device),
        }
    }

    fn train_state(&self, splats: &Splats<B>, lr_mean: f64) -> BrushTrainState {
        let c = &self.config;
        let p = |t: &Tensor<Inner, 2>| dptr(&t.clone().into_primitive().tensor()) as *mut f32;
        BrushTrainState {
            means: p(&splats.means.val().inner()),
            log_scales: p(&splats.log_scales.val().inner()),
            rotation: p(&splats.rotation.val().inner()),
            sh_coeffs: dptr(&splats.sh_coeffs.val().inner().into_primitive().tensor()) as *mut f32,
            raw_opacity: dptr(&splats.raw_opacity.val().inner().into_primitive().tensor()) as *mut f32,
            m_means: p(&self.mom_means.m),
            v_means: p(&self.mom_means.v),
            m_log_scales: p(&self.mom_scales.m),
            v_log_scales: p(&self.mom_scales.v),
            m_rotation: p(&self.mom_rot.m),
            v_rotation: p(&self.mom_rot.v),
            m_sh_coeffs: dptr(&self.mom_sh.m.clone().into_primitive().tensor()) as *mut f32,
            v_sh_coeffs: dptr(&self.mom_sh.v.clone().into_primitive().tensor()) as *mut f32,
            m_raw_opacity: dptr(&self.mom_opac.m.clone().into_primitive().tensor()) as *mut f32,
            v_raw_opacity: dptr(&self.mom_opac.v.clone().into_primitive().tensor()) as *mut f32,
            lr_mean: lr_mean as f32,
            lr_scale: c.lr_scale as f32,
            lr_rotation: c.lr_rotation as f32,
            lr_coeffs_dc: c.lr_coeffs_dc as f32,
            lr_opac: c.lr_opac as f32,
            // the reference steps every coefficient with lr_coeffs and lerps the higher bands back by
            // 1 / lr_coeffs_sh_scale (train.rs:336-349: "works as long as the update rule is linear"); the kernel
            // applies the scaled learning rate to bands >= 1 directly
            sh_rest_lr_scale: (1.0 / c.lr_coeffs_sh_scale) as f32,
        }
    }

    /// == SplatTrainer::step, train.rs:211-393, for the shape the reference asserts (one view, train.rs:216-219).
    /// Everything is enqueued on cubecl's stream; nothing synchronises except a refine.
    pub async fn step(
        &mut self,
        batch: SceneBatch<B>,
        splats: Splats<B>,
    ) -> Result<(Splats<B>, TrainStepStats<B>), anyhow::Error> {
        assert!(batch.gt_views.len() == 1, "Bigger batches aren't yet supported"); // train.rs:216-219
        let [_, img_h, img_w, channels] = batch.gt_images.dims();
        let device = splats.means.device();
        let n = splats.num_splats() as u32;
        let k = splats.sh_coeffs.dims()[1] as u32;
        let ctx_guard = ctx_for(&device).lock().unwrap();
        let ctx: &Ctx = &ctx_guard;

        // train.rs:229-237: the forward goes through the same Backend::render_splats as everywhere else (no autodiff
        // node is needed: the backward below is called directly)
        let camera = &batch.gt_views[0].camera;
        let (pred_image, aux): (Tensor<Inner, 3>, RenderAux<Inner>) =
            splats.valid().render(camera, glam::uvec2(img_w as u32, img_h as u32), false);
        let pred_prim = pred_image.clone().into_primitive().tensor();
        let stream: *mut c_void = stream_of(&pred_prim);

        // train.rs:249-268: loss = (1 - w) * L1 - w * SSIM and d loss / d pred, two launches
        let v_out = self.v_out.get_or_insert_with(|| Tensor::empty([img_h, img_w, 4], &device)).clone();
        let gt = batch.gt_images.clone().inner().reshape([img_h, img_w, channels]);
        unsafe {
            ctx.image_loss_backward(
                dptr(&pred_prim) as *const f32,
                dptr(&gt.into_primitive().tensor()) as *const f32,
                img_h as u32,
                img_w as u32,
                channels as u32,
                self.config.ssim_weight,
                self.config.ssim_window_size as u32,
                dptr(&v_out.clone().into_primitive().tensor()) as *mut f32,
                dptr(&self.loss.clone().into_primitive().tensor()) as *mut f32,
                stream,
            )?;
        }

        // train.rs:276-282
        let lr_mean = self.sched_mean.step() * batch.scene_extent;
        // train.rs:286: the statistics are accumulated once the warm-up is over
        let with_stats = self.iter > self.config.warmup_steps;
        let stat = |t: &Tensor<Inner, 1>| {
            if with_stats { dptr(&t.clone().into_primitive().tensor()) as *mut f32 } else { std::ptr::null_mut() }
        };

        // train.rs:361-365 (iter = the value before this step's increment); refine needs the pre-step parameters
        let do_refine = self.iter < self.config.max_refine_step
            && self.iter >= self.config.warmup_steps
            && self.iter % self.config.refine_every == 1;
        let pre_step = if do_refine { Some(splats.clone().detach_copy()) } else { None };

        // train.rs:270 (`loss.backward()`) + :284-316 (housekeeping) + :318-359 (optimizer) as rasterize_backwards and
        // ONE per-splat kernel; the parameters are updated in place
        self.opt_step += 1;
        let state = self.train_state(&splats, lr_mean);
        let c_aux: BrushAux = aux_of(&aux);
        unsafe {
            ctx.quat_normalize(
                state.rotation,
                n,
                dptr(&self.norm_rot.clone().into_primitive().tensor()) as *mut f32,
                stream,
            )?;
            ctx.render_backward_adam(
                n,
                k,
                &state,
                dptr(&self.norm_rot.clone().into_primitive().tensor()) as *const f32,
                dptr(&pred_prim) as *const f32,
                dptr(&v_out.into_primitive().tensor()) as *const f32,
                &c_aux,
                dptr(&self.v_xy.clone().into_primitive().tensor()) as *mut f32,
                stat(&self.grad_2d_accum),
                stat(&self.xy_grad_counts),
                AdamHyper { step: self.opt_step, ..AdamHyper::default() },
                stream,
            )?;
        }
        drop(ctx_guard);

        let mut refine_stats = None;
        let splats = if let Some(pre) = pre_step {
            let (s, r) = self.refine_splats(pre, splats).await?;
            refine_stats = Some(r);
            s
        } else {
            splats
        };
        self.iter += 1;

        let stats = TrainStepStats {
            pred_images: Tensor::from_inner(pred_image.unsqueeze()),
            gt_images: batch.gt_images,
            gt_views: batch.gt_views,
            auxes: vec![aux.into_autodiff()],
            loss: Tensor::from_inner(self.loss.clone()),
            lr_mean,
            lr_rotation: self.config.lr_rotation,
            lr_scale: self.config.lr_scale,
            lr_coeffs: self.config.lr_coeffs_dc,
            lr_opac: self.config.lr_opac,
            refine: refine_stats,
        };
        Ok((splats, stats))
    }

    /// == refine_splats, train.rs:395-553: classify + prune tests, block scan, compact-and-append in three launches.
    async fn refine_splats(
        &mut self,
        pre_step: Splats<B>,
        post_step: Splats<B>,
    ) -> Result<(Splats<B>, RefineStats), anyhow::Error> {
        let device = post_step.means.device();
        let n = post_step.num_splats() as u32;
        let k = post_step.sh_coeffs.dims()[1];
        let ctx_guard = ctx_for(&device).lock().unwrap();
        let ctx: &Ctx = &ctx_guard;
        let params = |s: &Splats<B>| BrushSplatParams {
            means: dptr(&s.means.val().inner().into_primitive().tensor()) as *mut f32,
            rotation: dptr(&s.rotation.val().inner().into_primitive().tensor()) as *mut f32,
            sh_coeffs: dptr(&s.sh_coeffs.val().inner().into_primitive().tensor()) as *mut f32,
            raw_opacity: dptr(&s.raw_opacity.val().inner().into_primitive().tensor()) as *mut f32,
            log_scales: dptr(&s.log_scales.val().inner().into_primitive().tensor()) as *mut f32,
        };
        let c = &self.config;
        let refine_step = self.iter / c.refine_every;
        let cfg = BrushRefineConfig {
            densify_grad_thresh: c.densify_grad_thresh,
            densify_size_thresh: c.densify_size_thresh,
            cull_alpha_thresh: c.cull_alpha_thresh,
            cull_scale_thresh: c.cull_scale_thresh,
            reset_alpha: (refine_step % c.reset_alpha_every_refine == 0) as i32, // train.rs:559-562
            reset_alpha_value: inverse_sigmoid(c.reset_alpha_value),
        };
        let stream: *mut c_void = stream_of(&self.grad_2d_accum.clone().into_primitive().tensor());
        // awaits the stream, as the reference awaits its argwhere_async calls (train.rs:424,450,506,548)
        let counts = unsafe {
            ctx.refine_plan(
                n,
                dptr(&self.grad_2d_accum.clone().into_primitive().tensor()) as *const f32,
                dptr(&self.xy_grad_counts.clone().into_primitive().tensor()) as *const f32,
                &params(&pre_step),
                &params(&post_step),
                &cfg,
                stream,
            )?
        };
        let m = counts.num_out as usize;
        // train.rs:474: the split offsets are N(0, 0.5) draws, one vec3 per split splat
        let samples: Tensor<Inner, 2> =
            Tensor::random([counts.num_split as usize, 3], Distribution::Normal(0.0, 0.5), &device);
        let out = Splats::<B>::from_raw(
            Tensor::empty([m, 3], &device),
            Tensor::empty([m, k, 3], &device),
            Tensor::empty([m, 4], &device),
            Tensor::empty([m], &device),
            Tensor::empty([m, 3], &device),
        );
        unsafe {
            ctx.refine_apply(
                n,
                k as u32,
                &params(&pre_step),
                &params(&post_step),
                dptr(&samples.into_primitive().tensor()) as *const f32,
                &cfg,
                &params(&out),
                stream,
            )?;
        }
        drop(ctx_guard);
        // train.rs:564-568: statistics and optimizer state start over (the Adam time with them)
        *self = Self { iter: self.iter, sched_mean: self.sched_mean.clone(), ..Self::new(&out, &self.config) };
        let stats = RefineStats {
            num_split: counts.num_split as usize,
            num_cloned: counts.num_cloned as usize,
            num_transparent_pruned: counts.alpha_pruned as usize,
            num_scale_pruned: counts.scale_pruned as usize,
        };
        Ok((out, stats))
    }
}
