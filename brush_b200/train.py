"""Host-side mirror of the brush-train step surface (crates/brush-train/src/train.rs).

    TrainConfig        == train.rs:19-87 (the fields the step uses)
    SceneBatch         == train.rs:89-94
    SplatTrainer.step  == train.rs:211-393 (render -> L1 loss -> backward -> xy-grad stats -> Adam)

Scope: the graded path is render fwd+bwd; loss and Adam here are single fused CUDA kernels in
libbrush_b200.so so that a benchmarked step is a complete training step.  SSIM (ssim_weight) and
refine_splats are not built yet: ssim_weight must be 0 and refinement is a no-op.
Multi-GPU: batch views are sharded over ranks; the dense per-Gaussian gradients are summed
with ONE NCCL all-reduce over a flat buffer before the (replicated) Adam step."""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass, field
from typing import List, Optional

import torch

from . import _lib
from .render import (Camera, RenderAux, Splats, _stream_ptr, get_context, make_uniforms, render_backward,
                     render_forward)


@dataclass
class TrainConfig:  # train.rs:19-87 defaults
    lr_mean: float = 1.6e-4
    lr_coeffs_dc: float = 0.004
    lr_coeffs_sh_scale: float = 20.0
    lr_opac: float = 0.05
    lr_scale: float = 0.01
    lr_rotation: float = 0.002
    ssim_weight: float = 0.0  # reference default 0.2; SSIM is a "next" row
    warmup_steps: int = 500
    refine_every: int = 100
    max_refine_step: int = 15000
    adam_beta1: float = 0.9
    adam_beta2: float = 0.999
    adam_eps: float = 1e-15  # train.rs:184


@dataclass
class SceneBatch:  # train.rs:89-94
    gt_images: torch.Tensor          # [B,H,W,C] f32 on the device, C = 3 or 4
    gt_views: List[Camera]
    scene_extent: float = 1.0


@dataclass
class TrainStepStats:  # train.rs:96-110
    pred_images: List[torch.Tensor]
    auxes: List[RenderAux]
    loss: torch.Tensor
    lr_mean: float
    lr_rotation: float
    lr_scale: float
    lr_coeffs: float
    lr_opac: float
    refine: Optional[object] = None


class SplatTrainer:
    """SplatTrainer (train.rs:151-198).  One instance per rank; `process_group` switches on the view-parallel
    gradient all-reduce."""

    # flat gradient layout [N*(13+3K)]: means 3 | xy 2 | log_scales 3 | quats 4 | sh 3K | opac 1
    def __init__(self, num_points: int, config: TrainConfig, device, sh_coeffs_per_channel: int = 1,
                 process_group=None, isect_capacity: Optional[int] = None):
        if config.ssim_weight != 0.0:
            raise NotImplementedError("SSIM loss is not built yet (SURVEY.md 8f-1); use ssim_weight=0")
        self.config = config
        self.device = torch.device(device)
        self.iter = 0
        self.pg = process_group
        self.isect_capacity = isect_capacity
        self._alloc(num_points, sh_coeffs_per_channel)

    def _alloc(self, n: int, k: int):
        f = dict(dtype=torch.float32, device=self.device)
        self.n, self.k = n, k
        per = 13 + 3 * k
        self.flat_grad = torch.zeros(n * per + 2 * n, **f)  # + the two refine statistics
        sizes = [3 * n, 2 * n, 3 * n, 4 * n, 3 * k * n, n, n, n]
        views = torch.split(self.flat_grad, sizes)
        self.g_means, self.g_xy, self.g_scales, self.g_quats = (views[0].view(n, 3), views[1].view(n, 2),
                                                                views[2].view(n, 3), views[3].view(n, 4))
        self.g_sh, self.g_opac = views[4].view(n, k, 3), views[5]
        self.stat_grad_norm, self.stat_visible = views[6], views[7]
        self.tmp = [torch.empty_like(v) for v in (self.g_means, self.g_xy, self.g_scales, self.g_quats, self.g_sh,
                                                  self.g_opac)]
        self.grad_2d_accum = torch.zeros(n, **f)   # train.rs:195-196
        self.xy_grad_counts = torch.zeros(n, **f)
        self.state = {}
        self.loss_buf = torch.zeros(1, **f)
        self.v_out = None

    def _adam_state(self, name, p):
        if name not in self.state or self.state[name][0].shape != p.shape:
            self.state[name] = (torch.zeros_like(p), torch.zeros_like(p))
        return self.state[name]

    def step(self, batch: SceneBatch, splats: Splats):
        cfg = self.config
        ctx = get_context(self.device)
        B = len(batch.gt_views)
        _, H, W, Cc = batch.gt_images.shape
        if self.v_out is None or self.v_out.shape[:2] != (H, W):
            self.v_out = torch.empty((H, W, 4), dtype=torch.float32, device=self.device)
        world = self.pg.size() if self.pg is not None else 1
        quats = splats.rotation.detach()
        qn = torch.sqrt(torch.sum(quats * quats, dim=1, keepdim=True))
        norm_rot = quats / qn  # gaussian_splats.rs:174-175
        p_means, p_scales = splats.means.detach(), splats.log_scales.detach()
        p_sh, p_opac = splats.sh_coeffs.detach(), splats.raw_opacity.detach()
        preds, auxes = [], []
        loss_total = torch.zeros(1, dtype=torch.float32, device=self.device)
        outs = (self.g_means, self.g_xy, self.g_scales, self.g_quats, self.g_sh, self.g_opac)
        for i, cam in enumerate(batch.gt_views):  # train.rs:229-237 (batch looped)
            u = make_uniforms(cam.world_to_local(), cam.focal((W, H)), (W, H), cam.center((W, H)))
            pred, aux = render_forward(u, p_means, p_scales, norm_rot, p_sh, p_opac, False,
                                       isect_capacity=self.isect_capacity, ctx=ctx)
            gt = batch.gt_images[i].contiguous()
            with torch.cuda.device(self.device):
                ctx.check(_lib.lib().brush_l1_loss_backward(
                    ctx.handle, pred.data_ptr(), gt.data_ptr(), H, W, Cc, self.v_out.data_ptr(),
                    self.loss_buf.data_ptr(), C.c_void_p(_stream_ptr(self.device))))
            loss_total += self.loss_buf
            dst = outs if i == 0 else self.tmp
            render_backward(aux, p_means, p_scales, norm_rot, p_sh.shape, p_opac, pred, self.v_out, ctx=ctx, out=dst)
            if i > 0:
                for a, b in zip(outs, self.tmp):
                    a += b
            # refine statistics are per view, BEFORE the reduction (the norm is nonlinear; train.rs:300-314)
            gx = dst[1]
            norm = torch.sqrt((gx[:, 0] * (W / 2.0)) ** 2 + (gx[:, 1] * (H / 2.0)) ** 2)
            if i == 0:
                self.stat_grad_norm.copy_(norm)
                self.stat_visible.zero_()
            else:
                self.stat_grad_norm.add_(norm)
            vis = (torch.arange(self.n, device=self.device) < aux.num_visible).float()
            self.stat_visible.index_add_(0, aux.global_from_compact_gid.long(), vis)
            preds.append(pred)
            auxes.append(aux)
        if self.pg is not None and world > 1:
            torch.distributed.all_reduce(self.flat_grad, group=self.pg)  # NCCL over NVLink
            torch.distributed.all_reduce(loss_total, group=self.pg)
        grad_scale = 1.0 / (B * world)  # loss is a mean over the global batch (train.rs:255)
        loss = loss_total * grad_scale
        # the quaternion normalisation's VJP (done by Burn autodiff in the reference)
        gq = self.g_quats
        gq.copy_((gq - norm_rot * torch.sum(norm_rot * gq, dim=1, keepdim=True)) / qn)
        if self.iter > cfg.warmup_steps:  # train.rs:286-316
            self.grad_2d_accum += self.stat_grad_norm * grad_scale
            self.xy_grad_counts = torch.where(self.stat_visible > 0, torch.ones_like(self.xy_grad_counts),
                                              self.xy_grad_counts)
        lr_mean = cfg.lr_mean * batch.scene_extent
        self.iter += 1
        self._adam(ctx, splats, lr_mean, grad_scale)
        stats = TrainStepStats(preds, auxes, loss, lr_mean, cfg.lr_rotation, cfg.lr_scale, cfg.lr_coeffs_dc,
                               cfg.lr_opac)
        return splats, stats

    def _adam(self, ctx, splats: Splats, lr_mean: float, grad_scale: float):
        cfg = self.config
        groups = (_lib.BrushAdamGroup * 5)()
        spec = [("means", splats.means, self.g_means, lr_mean, 0), ("opac", splats.raw_opacity, self.g_opac, cfg.lr_opac, 0),
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
        with torch.cuda.device(self.device):
            ctx.check(_lib.lib().brush_adam_step(ctx.handle, groups, 5, cfg.adam_beta1, cfg.adam_beta2, cfg.adam_eps,
                                                 self.iter, grad_scale, C.c_void_p(_stream_ptr(self.device))))
