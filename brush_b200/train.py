"""Host-side mirror of the brush-train step surface (crates/brush-train/src/train.rs).

    TrainConfig        == train.rs:19-87 (the fields the step uses)
    SceneBatch         == train.rs:89-94
    SplatTrainer.step  == train.rs:211-393 (render -> image loss -> backward -> xy-grad stats -> Adam -> refine)
    refine_splats      == train.rs:395-553

Scope: the graded path is render fwd+bwd; loss (L1 + SSIM), Adam, the rotation normalisation and
the refine statistics are CUDA kernels in libbrush_b200.so, so a benchmarked step is a complete training step with
no PyTorch arithmetic in it.  refine_splats runs once per `refine_every` steps and is host-orchestrated tensor
selection / concatenation, like the reference's (Burn ops + argwhere readbacks).
Multi-GPU: batch views are sharded over ranks; the dense per-Gaussian gradients (+ the two refine
statistics) are exchanged either by ONE brush_dp_adam_step launch per rank over NVLink peer memory (exchange =
"peer": sum + Adam + parameter broadcast fused) or by ONE NCCL all-reduce followed by the replicated Adam."""
from __future__ import annotations

import ctypes as C
import math
from dataclasses import dataclass
from typing import List, Optional

import torch

from . import _lib
from .parallel import ViewBuffers, FlatGradLayout, FlatParamLayout, PeerBuffers, allreduce_flat, global_grad_scale
from .render import (Camera, RenderAux, Splats, _stream_ptr, get_context, make_uniforms, poison_, poison_enabled,
                     render_backward, render_forward)


@dataclass
class TrainConfig:  # train.rs:19-87 defaults
    lr_mean: float = 1.6e-4            # ExponentialLrSchedulerConfig::new(lr_max, decay): initial lr ...
    lr_mean_decay: float = 1.0         # ... and per-step gamma (load_data.rs:51-67: 1e-2^(1/total_steps))
    lr_coeffs_dc: float = 0.004
    lr_coeffs_sh_scale: float = 20.0
    lr_opac: float = 0.05
    lr_scale: float = 0.01
    lr_rotation: float = 0.002
    ssim_weight: float = 0.2           # train.rs:56-57
    ssim_window_size: int = 11         # train.rs:59-60 (the kernel implements the default window only)
    scale_mean_lr_by_extent: bool = True
    warmup_steps: int = 500            # train.rs:22-31
    refine_every: int = 100
    max_refine_step: int = 15000
    reset_alpha_value: float = 0.004   # train.rs:33-54
    cull_alpha_thresh: float = 0.005
    cull_scale_thresh: float = 5.0
    reset_alpha_every_refine: int = 30
    densify_grad_thresh: float = 0.0002
    densify_size_thresh: float = 0.005
    seed: int = 42
    adam_beta1: float = 0.9
    adam_beta2: float = 0.999
    adam_eps: float = 1e-15  # train.rs:184


@dataclass
class SceneBatch:  # train.rs:89-94
    gt_images: torch.Tensor          # [B,H,W,C] f32, C = 3 or 4; on the device, or in pinned host memory
    gt_views: List[Camera]
    scene_extent: float = 1.0
    ready: Optional[object] = None     # CUDA event: the images' host->device copy (SplatTrainer.prefetch)
    slot: Optional[int] = None         # staging slot the images live in


@dataclass
class RefineStats:  # train.rs:112-118
    num_split: int
    num_cloned: int
    num_transparent_pruned: int
    num_scale_pruned: int


@dataclass
class TrainStepStats:  # train.rs:96-110
    pred_images: List[torch.Tensor]
    auxes: List[RenderAux]
    loss: torch.Tensor               # f32[1] on the device (mean over the global batch)
    lr_mean: float
    lr_rotation: float
    lr_scale: float
    lr_coeffs: float
    lr_opac: float
    refine: Optional[object] = None


class SplatTrainer:
    """SplatTrainer (train.rs:151-198).  One instance per rank; `process_group` switches on the view-parallel
    gradient all-reduce.  Steady state allocates nothing."""

    # flat buffer layout: means 3N | xy 2N | log_scales 3N | quats 4N | sh 3KN | opac N | grad-norm N | visible N
    def __init__(self, num_points: int, config: TrainConfig, device, sh_coeffs_per_channel: int = 1,
                 process_group=None, isect_capacity: Optional[int] = None, optimize: bool = True,
                 exchange: str = "nccl", fused_tail: bool = True):
        if config.ssim_weight > 0.0 and config.ssim_window_size != 11:
            raise ValueError("ssim_window_size must be 11 (the only window the loss kernel implements)")
        self.config = config
        self.device = torch.device(device)
        self.iter = 0
        self.pg = process_group
        self.isect_capacity = isect_capacity
        self.optimize = optimize
        self.fused_tail = fused_tail  # single device, one view per step: backward tail + Adam in one kernel
        if exchange not in ("nccl", "peer", "views", "hybrid"):
            raise ValueError("exchange must be 'nccl', 'peer', 'views' or 'hybrid'")
        # "peer": gradient sum + Adam + parameter broadcast as one launch over NVLink peer memory (parallel.py)
        # "views": no gradient exchange at all: 48-byte per-splat rows are published, every rank redoes the
        #          per-splat backward of every view and updates its own replica (one view per rank per step)
        # "hybrid": SH as 16-byte colour rows rebuilt on every rank (replicated SH Adam), the 11 small floats per splat
        #          through the "peer" kernel (one view per rank per step)
        multi = process_group is not None and process_group.size() > 1
        self.exchange = exchange if (multi or exchange == "views") else "nccl"
        self.peer: Optional[PeerBuffers] = None
        self.views: Optional[ViewBuffers] = None
        self.colour: Optional[ViewBuffers] = None
        self.colour_epoch = 0
        self.epoch = 0
        self._copy_stream = None
        # overlapped peer exchange (see _peer_step); BRUSH_DP_OVERLAP=0 restores the single-launch exchange
        import os

        self.overlap_exchange = os.environ.get("BRUSH_DP_OVERLAP", "1") != "0"
        self.side_ctas = int(os.environ.get("BRUSH_DP_SIDE_CTAS", "2"))
        self._xchg_stream, self._sh_pending = None, False
        self._alloc(num_points, sh_coeffs_per_channel)

    def _alloc(self, n: int, k: int):
        f = dict(dtype=torch.float32, device=self.device)
        self.n, self.k = n, k
        # Adam's bias-correction time: restarts with the optimizer state (train.rs:568 re-creates the optimizer in
        # refine_splats, so Burn's per-parameter `time` is 1 again on the first step after a refine)
        self.opt_step = 0
        self.layout = FlatGradLayout(n, k)
        self.param_layout = FlatParamLayout(n, k)
        if self.exchange in ("peer", "hybrid"):
            self.peer = PeerBuffers(self.layout, self.param_layout, self.pg, self.device)
            self.flat_grad = self.peer.grad
            self.peer_m = torch.zeros(self.param_layout.total, **f)
            self.peer_v = torch.zeros(self.param_layout.total, **f)
        else:
            self.flat_grad = self.layout.alloc(self.device)
        if self.exchange == "views":
            self.views = ViewBuffers(n, self.pg, self.device)
        if self.exchange == "hybrid":
            self.colour = ViewBuffers(n, self.pg, self.device, stride=4)
        (self.g_means, self.g_xy, self.g_scales, self.g_quats, self.g_sh, self.g_opac, self.stat_grad_norm,
         self.stat_visible, self.g_loss) = self.layout.views(self.flat_grad)
        self.tmp = None  # second gradient set, only for batches of more than one view per rank
        self.grad_2d_accum = torch.zeros(n, **f)   # train.rs:195-196
        self.xy_grad_counts = torch.zeros(n, **f)
        self.norm_rot = torch.empty((n, 4), **f)
        if poison_enabled():
            poison_(self.norm_rot)
        self.state = {}
        # two sets of loss scalars, alternating by step: the loss a step returns stays valid while the NEXT step
        # runs, so a caller may read it back from another stream, one step behind, without a copy on this one
        self._loss_sets = [(torch.zeros(1, **f), torch.zeros(1, **f)) for _ in range(2)]
        self.loss_total, self.loss_buf = self._loss_sets[0]
        self.v_out = None
        self.gt_dev = None
        self.render_bufs: List = []

    def prefetch(self, batch: SceneBatch) -> SceneBatch:
        """Starts the host->device copy of a (pinned) batch on a side stream and returns the batch with device
        images; step() waits for the copy on its own stream.  Two staging slots alternate, so the copy of step
        t+1 runs under the kernels of step t -- what the reference's SceneLoader prefetch does for its uploads
        (crates/brush-dataset/src/scene_loader.rs:16-19)."""
        if batch.gt_images.is_cuda:
            return batch
        dev = self.device
        if self._copy_stream is None:
            self._copy_stream = torch.cuda.Stream(dev)
            self._slots, self._slot_free, self._next_slot = [None, None], [None, None], 0
            self._slots_u8 = [None, None]
        k = self._next_slot
        self._next_slot ^= 1
        if self._slots[k] is None or self._slots[k].shape != batch.gt_images.shape:
            self._slots[k] = torch.empty(batch.gt_images.shape, dtype=torch.float32, device=dev)
        quantized = batch.gt_images.dtype == torch.uint8  # bytes cross PCIe, /255 happens on the device
        if quantized and (self._slots_u8[k] is None or self._slots_u8[k].shape != batch.gt_images.shape):
            self._slots_u8[k] = torch.empty(batch.gt_images.shape, dtype=torch.uint8, device=dev)
        with torch.cuda.stream(self._copy_stream):
            if self._slot_free[k] is not None:
                self._copy_stream.wait_event(self._slot_free[k])  # the step that last read this slot is done
            if quantized:
                self._slots_u8[k].copy_(batch.gt_images, non_blocking=True)
                self._dequantize(self._slots_u8[k], self._slots[k], self._copy_stream.cuda_stream)
            else:
                self._slots[k].copy_(batch.gt_images, non_blocking=True)
            ready = torch.cuda.Event()
            ready.record(self._copy_stream)
        return SceneBatch(self._slots[k], batch.gt_views, batch.scene_extent, ready, k)

    def _dequantize(self, src_u8: torch.Tensor, dst_f32: torch.Tensor, stream_ptr: int) -> None:
        """u8 -> f32 / 255 on the device (== image_to_tensor of brush-dataset, after the upload)."""
        ctx = get_context(self.device)
        ctx.check(_lib.lib().brush_image_u8_to_f32(ctx.handle, src_u8.data_ptr(), src_u8.numel(), dst_f32.data_ptr(),
                                                   C.c_void_p(stream_ptr)))

    def adopt(self, splats: Splats) -> Splats:
        """Peer mode: moves the parameters into this rank's peer-mapped replica and returns Splats whose tensors
        are views of it (brush_dp_adam_step updates every rank's replica in place).  A no-op otherwise."""
        if self.peer is None:
            return splats
        pm, po, psh, pr, ps, _, _ = self.param_layout.views(self.peer.param)
        for dst, src in ((pm, splats.means), (po, splats.raw_opacity), (psh, splats.sh_coeffs), (pr, splats.rotation),
                         (ps, splats.log_scales)):
            dst.copy_(src.detach())
        torch.cuda.synchronize(self.device)
        torch.distributed.barrier(self.pg)
        return Splats.from_views(pm, psh, pr, po, ps)

    def _adam_state(self, name, p):
        if name not in self.state or self.state[name][0].shape != p.shape:
            self.state[name] = (torch.zeros_like(p), torch.zeros_like(p))
        return self.state[name]

    def step(self, batch: SceneBatch, splats: Splats):
        """train.rs:211-393.  Everything is enqueued on the current stream; nothing synchronises."""
        cfg, dev, L = self.config, self.device, _lib.lib()
        ctx = get_context(dev)
        st = C.c_void_p(_stream_ptr(dev))
        self.loss_total, self.loss_buf = self._loss_sets[self.iter & 1]
        B = len(batch.gt_views)
        _, H, W, Cc = batch.gt_images.shape
        if self.v_out is None or tuple(self.v_out.shape[:2]) != (H, W):
            self.v_out = torch.empty((H, W, 4), dtype=torch.float32, device=dev)
            if poison_enabled():
                poison_(self.v_out)
            self.render_bufs = []
        world = self.pg.size() if self.pg is not None else 1
        p_means, p_scales = splats.means.detach(), splats.log_scales.detach()
        p_sh, p_opac, p_rot = splats.sh_coeffs.detach(), splats.raw_opacity.detach(), splats.rotation.detach()
        outs = (self.g_means, self.g_xy, self.g_scales, self.g_quats, self.g_sh, self.g_opac)
        preds, auxes = [], []
        views = self.views is not None and self.optimize and B == 1
        hybrid = self.colour is not None
        if hybrid and (B != 1 or not self.optimize):
            raise ValueError("exchange='hybrid' takes one view per rank per step and needs the optimizer")
        fused = self.fused_tail and self.optimize and world == 1 and B == 1 and self.peer is None and not views
        # train.rs:361-365 (iter = the value before this step's increment)
        do_refine = (self.iter < cfg.max_refine_step and self.iter >= cfg.warmup_steps
                     and self.iter % cfg.refine_every == 1 and self.optimize)
        with torch.cuda.device(dev):
            if self._sh_pending:
                if do_refine or B != 1:
                    self.finish()  # this step reads sh_coeffs before its (first) forward gets to the colours
                else:
                    ctx.set_sh_ready_event(self._ev_sh.cuda_event)  # the forward waits just before it needs colours
                    self._sh_pending = False
            pre_step = [p.detach().clone() for p in (splats.means, splats.sh_coeffs, splats.rotation,
                                                     splats.raw_opacity, splats.log_scales)] if do_refine else None
            if batch.ready is not None:
                torch.cuda.current_stream(dev).wait_event(batch.ready)
            ctx.check(L.brush_quat_normalize(ctx.handle, p_rot.data_ptr(), self.n, self.norm_rot.data_ptr(), st))
            for i, cam in enumerate(batch.gt_views):  # train.rs:229-237 (batch looped)
                gt = batch.gt_images[i]
                if not gt.is_cuda or gt.dtype == torch.uint8:
                    # the view feed hands over pinned host images and / or bytes: H2D and /255 on the compute stream
                    if self.gt_dev is None or self.gt_dev.shape != gt.shape:
                        self.gt_dev = torch.empty(gt.shape, dtype=torch.float32, device=dev)
                    if gt.dtype == torch.uint8:
                        gt_u8 = gt if gt.is_cuda else gt.to(dev, non_blocking=True)
                        self._dequantize(gt_u8, self.gt_dev, _stream_ptr(dev))
                    else:
                        self.gt_dev.copy_(gt, non_blocking=True)
                    gt = self.gt_dev
                u = make_uniforms(cam.world_to_local(), cam.focal((W, H)), (W, H), cam.center((W, H)))
                reuse = self.render_bufs[i] if i < len(self.render_bufs) else None
                pred, aux = render_forward(u, p_means, p_scales, self.norm_rot, p_sh, p_opac, False,
                                           isect_capacity=self.isect_capacity, ctx=ctx, reuse=reuse)
                if reuse is None:
                    self.render_bufs.append((pred, aux))
                # loss = (1-w) L1 - w SSIM and d loss / d pred in two launches (train.rs:249-268, ssim.rs)
                ctx.check(L.brush_image_loss_backward(ctx.handle, pred.data_ptr(), gt.data_ptr(), H, W, Cc,
                                                      cfg.ssim_weight, cfg.ssim_window_size, self.v_out.data_ptr(),
                                                      (self.loss_total if i == 0 else self.loss_buf).data_ptr(), st))
                if i == 0:
                    dst = outs
                else:
                    self.loss_total.add_(self.loss_buf)
                    if self.tmp is None:
                        self.tmp = tuple(torch.empty_like(v) for v in outs)
                    dst = self.tmp
                if views:
                    # publish this view's per-splat rows; the dense gradients are never formed (see _views_step)
                    self.epoch += 1
                    vx = self.views.exchange(self.epoch)
                    caux = aux._c()
                    ctx.check(L.brush_render_backward_rows(ctx.handle, self.n, self.k, pred.data_ptr(),
                                                           self.v_out.data_ptr(), C.byref(caux),
                                                           self.loss_total.data_ptr(), C.byref(vx), st))
                    preds.append(pred)
                    auxes.append(aux)
                    break
                if fused:
                    # single device, one view: per-splat backward + normalisation VJP + Adam + refine statistics in
                    # one kernel; the dense gradients are never materialised
                    lr_mean = self._lr_mean(batch)
                    self._backward_adam(ctx, splats, aux, pred, lr_mean, self.iter > cfg.warmup_steps, st)
                    preds.append(pred)
                    auxes.append(aux)
                    break
                elif hybrid:
                    # small gradients densely + this view's colour rows (16 B per splat) published to every rank
                    self.colour_epoch += 1
                    cx = self.colour.exchange(self.colour_epoch)
                    caux = aux._c()
                    ctx.check(L.brush_render_backward_hybrid(
                        ctx.handle, self.n, self.k, p_means.data_ptr(), p_scales.data_ptr(), self.norm_rot.data_ptr(),
                        p_opac.data_ptr(), pred.data_ptr(), self.v_out.data_ptr(), C.byref(caux), dst[0].data_ptr(),
                        dst[1].data_ptr(), dst[2].data_ptr(), dst[3].data_ptr(), dst[5].data_ptr(),
                        self.loss_total.data_ptr(), C.byref(cx), st))
                else:
                    render_backward(aux, p_means, p_scales, self.norm_rot, p_sh.shape, p_opac, pred, self.v_out,
                                    ctx=ctx, out=dst)
                # refine statistics are per view, BEFORE the reduction (the norm is nonlinear; train.rs:300-314)
                caux = aux._c()
                ctx.check(L.brush_xy_grad_stats(ctx.handle, dst[1].data_ptr(), C.byref(caux), self.n, W, H,
                                                int(i > 0), self.stat_grad_norm.data_ptr(),
                                                self.stat_visible.data_ptr(), st))
                if i > 0:
                    for a, b in zip(outs, self.tmp):
                        a.add_(b)
                preds.append(pred)
                auxes.append(aux)
            grad_scale = global_grad_scale(B, world)
            lr_mean = self._lr_mean(batch)
            if not fused and not views:
                # VJP of the rotation normalisation (Burn autodiff in the reference); linear: commutes with the sum
                ctx.check(L.brush_quat_normalize_vjp(ctx.handle, p_rot.data_ptr(), self.n, self.g_quats.data_ptr(),
                                                     st))
            self.iter += 1
            self.opt_step += 1
            if views:
                with_stats = self.iter - 1 > cfg.warmup_steps  # train.rs:286-316
                self._views_step(ctx, splats, vx, lr_mean, grad_scale, with_stats, st)
                loss = self.loss_buf  # grad_scale * sum over ranks, written by the step kernel
            elif self.peer is not None:
                if splats.means.data_ptr() != self.peer.param.data_ptr() + 4 * self.param_layout.offset("means"):
                    raise ValueError("peer exchange: pass the Splats returned by SplatTrainer.adopt()")
                self.g_loss[:1].copy_(self.loss_total)
                if hybrid:
                    # SH: rebuilt from every view's colour rows and stepped on this rank's replica; needs the means the
                    # forwards rendered with, so it runs BEFORE the small groups are updated
                    m_sh, v_sh = self._adam_state("sh", splats.sh_coeffs)
                    ctx.check(L.brush_sh_views_adam_step(
                        ctx.handle, self.n, self.k, p_means.data_ptr(), p_sh.data_ptr(), m_sh.data_ptr(), v_sh.data_ptr(),
                        cfg.lr_coeffs_dc, 1.0 / cfg.lr_coeffs_sh_scale, C.byref(cx), None, cfg.adam_beta1,
                        cfg.adam_beta2, cfg.adam_eps, self.opt_step, grad_scale, st))
                stat_norm, stat_vis = self._peer_step(ctx, lr_mean, grad_scale, st, skip_sh=hybrid)
                loss = self.loss_buf  # grad_scale * sum over ranks, written by the exchange kernel
            else:
                if self.pg is not None and world > 1:
                    allreduce_flat(self.flat_grad, self.pg)  # ONE NCCL all-reduce over NVLink
                    allreduce_flat(self.loss_total, self.pg)
                stat_norm, stat_vis = self.stat_grad_norm, self.stat_visible
                if self.optimize and not fused:
                    self._adam(ctx, splats, lr_mean, grad_scale, st)
                loss = self.loss_total * grad_scale if (B * world) > 1 else self.loss_total
            if self.iter - 1 > cfg.warmup_steps and not views and not fused:  # train.rs:286-316
                self.grad_2d_accum.add_(stat_norm, alpha=grad_scale)
                self.xy_grad_counts.add_(stat_vis)  # select_assign == scatter-ADD in Burn 0.16: a visibility count
            if batch.slot is not None:
                self._slot_free[batch.slot] = torch.cuda.Event()
                self._slot_free[batch.slot].record(torch.cuda.current_stream(dev))
            refine = None
            if do_refine:
                self.finish()
                loss = loss.clone()
                splats, refine = self.refine_splats(pre_step, splats)
        stats = TrainStepStats(preds, auxes, loss, lr_mean, cfg.lr_rotation, cfg.lr_scale, cfg.lr_coeffs_dc,
                               cfg.lr_opac, refine)
        return splats, stats

    def refine_splats(self, pre_step, post: Splats, samples: Optional[torch.Tensor] = None):
        """train.rs:395-553: clone small / split large high-gradient splats, prune transparent and huge ones, reset
        opacities every `reset_alpha_every_refine` refines, reset the statistics and the optimizer state.
        `pre_step` = (means, sh, rotation, raw_opacity, log_scales) BEFORE this step's optimizer update.
        As in the reference, the edits it makes to the SPLIT splats themselves (smaller scale, offset mean) are
        applied to clones that are then dropped (train.rs:401-402,481-507 vs :519): only the appended copies
        carry them.  Every rank takes the same decisions: the statistics are the reduced ones and the split
        offsets come from a generator seeded with (config.seed, iter).  `samples` [num_split, 3] overrides the
        N(0, 0.5) draws (tests).  The arithmetic runs in three kernels behind brush_refine_plan /
        brush_refine_apply (csrc/refine.cu); torch only allocates the new tensors and draws the samples."""
        cfg, dev, L = self.config, self.device, _lib.lib()
        ctx = get_context(dev)
        st = C.c_void_p(_stream_ptr(dev))
        f32 = lambda t: t.detach().contiguous().float()
        pre = [f32(t) for t in pre_step]  # means, sh, rotation, raw_opacity, log_scales
        post = [f32(t) for t in (post.means, post.sh_coeffs, post.rotation, post.raw_opacity, post.log_scales)]

        def params(ts):  # BrushSplatParams order: means, rotation, sh, raw_opacity, log_scales
            return _lib.BrushSplatParams(ts[0].data_ptr(), ts[2].data_ptr(), ts[1].data_ptr(), ts[3].data_ptr(),
                                         ts[4].data_ptr())

        refine_step = (self.iter - 1) // cfg.refine_every  # self.iter was already incremented
        rcfg = _lib.BrushRefineConfig(cfg.densify_grad_thresh, cfg.densify_size_thresh, cfg.cull_alpha_thresh,
                                      cfg.cull_scale_thresh, int(refine_step % cfg.reset_alpha_every_refine == 0),
                                      cfg.reset_alpha_value)
        counts = _lib.BrushRefineCounts()
        c_pre, c_post = params(pre), params(post)
        with torch.cuda.device(dev):
            # decisions + prune tests + counts on the device; synchronises like the reference's argwhere().await
            ctx.check(L.brush_refine_plan(ctx.handle, self.n, self.grad_2d_accum.data_ptr(),
                                          self.xy_grad_counts.data_ptr(), C.byref(c_pre), C.byref(c_post),
                                          C.byref(rcfg), C.byref(counts), st))
            # brush_refine_plan has just synchronised the stream: the cheap place to surface an intersection
            # overflow of any forward since the last refine (raises CapacityError; the reference truncates silently)
            ctx.sync_check(_stream_ptr(dev))
            n_split, n_out = int(counts.num_split), int(counts.num_out)
            if samples is None:
                gen = torch.Generator(device=dev)
                gen.manual_seed((cfg.seed << 20) + self.iter)
                samples = torch.randn((max(n_split, 1), 3), device=dev, generator=gen) * 0.5
            samples = f32(samples)
            if n_split and tuple(samples.shape) != (n_split, 3):
                raise ValueError(f"samples must be [{n_split}, 3]")
            k = post[1].shape[1]
            f = dict(dtype=torch.float32, device=dev)
            new = [torch.empty((n_out, 3), **f), torch.empty((n_out, k, 3), **f), torch.empty((n_out, 4), **f),
                   torch.empty((n_out,), **f), torch.empty((n_out, 3), **f)]
            c_out = params(new)
            ctx.check(L.brush_refine_apply(ctx.handle, self.n, k, C.byref(c_pre), C.byref(c_post), samples.data_ptr(),
                                           C.byref(rcfg), C.byref(c_out), st))
        n_clone, alpha_pruned, scale_pruned = int(counts.num_cloned), int(counts.alpha_pruned), int(counts.scale_pruned)
        new = [new[0], new[2], new[1], new[3], new[4]]  # -> means, rotation, sh, opacity, log_scales (as below)
        it = self.iter
        self._alloc(new[0].shape[0], self.k)  # reset_stats + a fresh optimizer (train.rs:550-554)
        self.iter = it
        out = self.adopt(Splats(new[0], new[2], new[1], new[3], new[4]))
        return out, RefineStats(n_split, n_clone, alpha_pruned, scale_pruned)

    def _peer_step(self, ctx, lr_mean: float, grad_scale: float, st, skip_sh: bool = False):
        """brush_dp_adam_step: sum of the ranks' gradients on the owned slice -> Adam (train.rs:318-359) -> new
        parameters into every replica; the refine statistics ride along as plain-sum segments.

        With `overlap_exchange` (default) the exchange is TWO launches: the small groups (means, opacity, rotation,
        scales: 11 floats per splat, plus statistics and loss) on the step's stream, and the SH group (3K floats per
        splat, 81 % of the bytes) on a side stream.  The next forward needs SH only for the colours, which it
        evaluates last (brush_ctx_set_sh_ready_event): projection, both sorts, scan, intersection lists and tile bins
        of step t+1 run while step t's SH exchange is still on the wire.  Parameters are exact, never stale."""
        cfg, pl, gl, peer = self.config, self.param_layout, self.layout, self.peer
        lrs = {"means": lr_mean, "opacity": cfg.lr_opac, "sh": cfg.lr_coeffs_dc, "rotation": cfg.lr_rotation,
               "log_scales": cfg.lr_scale}

        def seg_array(names):
            segs = (_lib.BrushDpSegment * len(names))()
            for i, name in enumerate(names):
                segs[i].src_offset = gl.offset(pl.GRAD_OF[name])
                segs[i].dst_offset = pl.offset(name)
                segs[i].count = pl.padded_size(name)
                segs[i].adam = int(name in lrs and self.optimize)
                segs[i].lr = lrs.get(name, 0.0)
                segs[i].sh_coeffs_per_channel = self.k if name == "sh" else 0
                segs[i].sh_rest_lr_scale = 1.0 / cfg.lr_coeffs_sh_scale
            return segs

        def launch(names, lane, ctas, loss_out, stream_ptr):
            sp = _lib.BrushDpStep()
            sp.world, sp.rank = peer.world, peer.rank
            for r in range(peer.world):
                sp.src[r], sp.dst[r], sp.flags[r] = peer.grad_ptrs[r], peer.param_ptrs[r], peer.flag_ptrs[r]
            sp.src_multicast, sp.dst_multicast = peer.grad_mc or None, peer.param_mc or None
            sp.m, sp.v = self.peer_m.data_ptr(), self.peer_v.data_ptr()
            sp.loss_offset, sp.loss_out = gl.offset("loss"), loss_out
            sp.epoch, sp.lane, sp.ctas_per_sm = self.epoch, lane, ctas
            segs = seg_array(names)
            ctx.check(_lib.lib().brush_dp_adam_step(ctx.handle, C.byref(sp), segs, len(segs), cfg.adam_beta1,
                                                    cfg.adam_beta2, cfg.adam_eps, self.opt_step, grad_scale,
                                                    C.c_void_p(stream_ptr)))

        names = [n for n in pl.NAMES if not (skip_sh and n == "sh")]  # hybrid: SH never goes through this exchange
        if not self.optimize:  # gradients only: leave the parameters alone, still reduce the statistics
            names = ["grad_norm", "visible"]
        self.epoch += 1
        main = torch.cuda.current_stream(self.device)
        if self.overlap_exchange and self.optimize and "sh" in names:
            if self._xchg_stream is None:
                self._xchg_stream = torch.cuda.Stream(self.device)
                self._ev_grads, self._ev_sh = torch.cuda.Event(), torch.cuda.Event()
            self._ev_grads.record(main)  # this rank's dense gradients are complete
            launch([n for n in names if n != "sh"], 0, 0, self.loss_buf.data_ptr(), main.cuda_stream)
            side = self._xchg_stream
            side.wait_event(self._ev_grads)
            launch(["sh"], 1, self.side_ctas, None, side.cuda_stream)
            self._ev_sh.record(side)
            self._sh_pending = True
        else:
            launch(names, 0, 0, self.loss_buf.data_ptr(), main.cuda_stream)
        views = pl.views(peer.param)
        return views[5], views[6]

    def finish(self):
        """Makes the current stream wait for an SH exchange that is still in flight on the side stream (overlapped
        peer exchange).  step() hands the wait to the next forward instead; call this before reading the parameters
        from any other code (evaluation renders go through the ctx and wait by themselves)."""
        if self._sh_pending:
            torch.cuda.current_stream(self.device).wait_event(self._ev_sh)
            get_context(self.device).set_sh_ready_event(None)
            self._sh_pending = False

    def _lr_mean(self, batch: SceneBatch) -> float:
        """ExponentialLrScheduler (Burn 0.16): step k (1-based) returns initial_lr * gamma^(k-1); train.rs:276-278."""
        cfg = self.config
        lr_mean = cfg.lr_mean * cfg.lr_mean_decay ** self.iter
        if cfg.scale_mean_lr_by_extent:
            lr_mean *= batch.scene_extent
        return lr_mean

    def _train_state(self, splats: Splats, lr_mean: float):
        cfg = self.config
        S = _lib.BrushTrainState()
        for field, name, p in (("means", "means", splats.means), ("raw_opacity", "opac", splats.raw_opacity),
                               ("sh_coeffs", "sh", splats.sh_coeffs), ("rotation", "rot", splats.rotation),
                               ("log_scales", "scale", splats.log_scales)):
            m, v = self._adam_state(name, p)
            setattr(S, field, p.data_ptr())
            setattr(S, "m_" + field, m.data_ptr())
            setattr(S, "v_" + field, v.data_ptr())
        S.lr_mean, S.lr_scale, S.lr_rotation = lr_mean, cfg.lr_scale, cfg.lr_rotation
        S.lr_coeffs_dc, S.lr_opac, S.sh_rest_lr_scale = cfg.lr_coeffs_dc, cfg.lr_opac, 1.0 / cfg.lr_coeffs_sh_scale
        return S

    def _views_step(self, ctx, splats: Splats, vx, lr_mean: float, grad_scale: float, with_stats: bool, st):
        """brush_views_adam_step: per-splat backward of every rank's view + Adam on this rank's replica."""
        cfg = self.config
        S = self._train_state(splats, lr_mean)
        accum = (self.grad_2d_accum.data_ptr(), self.xy_grad_counts.data_ptr()) if with_stats else (None, None)
        ctx.check(_lib.lib().brush_views_adam_step(
            ctx.handle, self.n, self.k, C.byref(S), self.norm_rot.data_ptr(), C.byref(vx), accum[0], accum[1],
            self.loss_buf.data_ptr(), cfg.adam_beta1, cfg.adam_beta2, cfg.adam_eps, self.opt_step, grad_scale, st))

    def _backward_adam(self, ctx, splats: Splats, aux, pred, lr_mean: float, with_stats: bool, st):
        cfg = self.config
        S = self._train_state(splats, lr_mean)
        caux = aux._c()
        accum = (self.grad_2d_accum.data_ptr(), self.xy_grad_counts.data_ptr()) if with_stats else (None, None)
        ctx.check(_lib.lib().brush_render_backward_adam(
            ctx.handle, self.n, self.k, C.byref(S), self.norm_rot.data_ptr(), pred.data_ptr(), self.v_out.data_ptr(),
            C.byref(caux), None, accum[0], accum[1], cfg.adam_beta1, cfg.adam_beta2, cfg.adam_eps, self.opt_step + 1, 1.0,
            st))

    def _adam(self, ctx, splats: Splats, lr_mean: float, grad_scale: float, st):
        cfg = self.config
        groups = (_lib.BrushAdamGroup * 5)()
        spec = [("means", splats.means, self.g_means, lr_mean, 0),
                ("opac", splats.raw_opacity, self.g_opac, cfg.lr_opac, 0),
                ("sh", splats.sh_coeffs, self.g_sh, cfg.lr_coeffs_dc, self.k),
                ("rot", splats.rotation, self.g_quats, cfg.lr_rotation, 0),
                ("scale", splats.log_scales, self.g_scales, cfg.lr_scale, 0)]  # order of train.rs:320-356
        for i, (name, p, g, lr, shk) in enumerate(spec):
            m, v = self._adam_state(name, p)
            groups[i].param, groups[i].grad = p.data_ptr(), g.data_ptr()
            groups[i].m, groups[i].v = m.data_ptr(), v.data_ptr()
            groups[i].count = p.numel()
            groups[i].lr = lr
            groups[i].sh_coeffs_per_channel = shk
            groups[i].sh_rest_lr_scale = 1.0 / cfg.lr_coeffs_sh_scale
        ctx.check(_lib.lib().brush_adam_step(ctx.handle, groups, 5, cfg.adam_beta1, cfg.adam_beta2, cfg.adam_eps,
                                             self.opt_step, grad_scale, st))
