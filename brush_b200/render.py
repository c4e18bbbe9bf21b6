"""Host-side mirror of the reference operator surface for the splat path, above the C ABI.

    Camera            == crates/brush-render/src/camera.rs:1-58
    RenderAux         == crates/brush-render/src/lib.rs:20-63
    render_splats     == brush_render::Backend::render_splats (lib.rs:66-86), both the plain backend
                         (render.rs:325-350) and the autodiff one (render.rs:366-463 + 465-626)
    Splats            == crates/brush-render/src/gaussian_splats.rs:24-224 (the parts on the path)

PyTorch is only plumbing here (device memory, streams, the autograd tape standing in for Burn's);
every kernel runs inside libbrush_b200.so.  There is no fallback path."""
from __future__ import annotations

import ctypes as C
import math
import os
from dataclasses import dataclass
from typing import Optional, Tuple

import numpy as np
import torch

from . import _lib
from ._lib import BrushAux, BrushUniforms, Context

TILE_WIDTH = 16
REFERENCE_MAX_INTERSECTS = 128 * 65535  # render.rs:204-206


def poison_enabled() -> bool:
    """BRUSH_POISON=1: every caller-allocated aux / output buffer is noise-filled before the library call, the way
    the reference's `create_tensor` does under test (crates/brush-kernel/src/lib.rs:134-147 fills "empty" buffers
    with -12345), so a kernel that relies on zero-initialised memory fails loudly."""
    return os.environ.get("BRUSH_POISON", "0") == "1"


def poison_(*tensors):
    for t in tensors:
        if t is None:
            continue
        if t.dtype.is_floating_point:
            t.fill_(-12345.0)
        else:
            t.view(torch.int32).fill_(-0x30303031)  # 0xCFCFCFCF


def fov_to_focal(fov_rad: float, pixels: int) -> float:  # camera.rs:52-54
    return 0.5 * float(pixels) / math.tan(fov_rad * 0.5)


def focal_to_fov(focal: float, pixels: int) -> float:  # camera.rs:57-59
    return 2.0 * math.atan(float(pixels) / (2.0 * focal))


class Camera:
    """camera.rs:1-50.  `rotation` is a quaternion in glam order (x, y, z, w); camera looks down +z."""

    def __init__(self, position, rotation, fov_x: float, fov_y: float, center_uv=(0.5, 0.5)):
        self.position = np.asarray(position, dtype=np.float64)
        self.rotation = np.asarray(rotation, dtype=np.float64)
        self.fov_x, self.fov_y = float(fov_x), float(fov_y)
        self.center_uv = (float(center_uv[0]), float(center_uv[1]))

    def focal(self, img_size) -> Tuple[float, float]:
        return (float(np.float32(fov_to_focal(self.fov_x, img_size[0]))),
                float(np.float32(fov_to_focal(self.fov_y, img_size[1]))))

    def center(self, img_size) -> Tuple[float, float]:
        return (float(np.float32(self.center_uv[0]) * np.float32(img_size[0])),
                float(np.float32(self.center_uv[1]) * np.float32(img_size[1])))

    def _rotmat(self) -> np.ndarray:
        x, y, z, w = self.rotation / np.linalg.norm(self.rotation)
        return np.array([[1 - 2 * (y * y + z * z), 2 * (x * y - w * z), 2 * (x * z + w * y)],
                         [2 * (x * y + w * z), 1 - 2 * (x * x + z * z), 2 * (y * z - w * x)],
                         [2 * (x * z - w * y), 2 * (y * z + w * x), 1 - 2 * (x * x + y * y)]], dtype=np.float64)

    def local_to_world(self) -> np.ndarray:
        m = np.eye(4, dtype=np.float64)
        m[:3, :3] = self._rotmat()
        m[:3, 3] = self.position
        return m

    def world_to_local(self) -> np.ndarray:
        """Rigid inverse (glam's general Mat4::inverse differs by <= 1 ulp; viewmat is an ABI input anyway)."""
        r = self._rotmat()
        m = np.eye(4, dtype=np.float64)
        m[:3, :3] = r.T
        m[:3, 3] = -r.T @ self.position
        return m.astype(np.float32)


def make_uniforms(viewmat: np.ndarray, focal, img_size, pixel_center) -> BrushUniforms:
    """render.rs:102-116.  viewmat[r, c] row-major numpy -> column-major mat4."""
    u = BrushUniforms()
    cols = np.asarray(viewmat, dtype=np.float32).T.reshape(-1)
    for i in range(16):
        u.viewmat[i] = float(cols[i])
    u.focal[0], u.focal[1] = float(np.float32(focal[0])), float(np.float32(focal[1]))
    u.img_size[0], u.img_size[1] = int(img_size[0]), int(img_size[1])
    u.tile_bounds[0] = (int(img_size[0]) + TILE_WIDTH - 1) // TILE_WIDTH
    u.tile_bounds[1] = (int(img_size[1]) + TILE_WIDTH - 1) // TILE_WIDTH
    u.pixel_center[0], u.pixel_center[1] = float(np.float32(pixel_center[0])), float(np.float32(pixel_center[1]))
    return u


def sh_degree_from_coeffs(coeffs_per_channel: int) -> int:  # render.rs:44-53
    table = {1: 0, 4: 1, 9: 2, 16: 3, 25: 4}
    if coeffs_per_channel not in table:
        raise ValueError(f"Invalid nr. of sh bases {coeffs_per_channel}")
    return table[coeffs_per_channel]


def sh_coeffs_for_degree(degree: int) -> int:  # render.rs:40-42
    return (degree + 1) ** 2


@dataclass
class RenderAux:
    """lib.rs:20-33.  All tensors live on the render device; int tensors are int32 like Burn's."""

    projected_splats: torch.Tensor       # f32[N,9]
    uniforms_buffer: torch.Tensor        # i32[28]
    num_intersections: torch.Tensor      # i32[1]
    num_visible: torch.Tensor            # i32[1]
    final_index: torch.Tensor            # i32[H,W]
    cum_tiles_hit: torch.Tensor          # i32[N]
    tile_bins: torch.Tensor              # i32[Ty,Tx,2]
    compact_gid_from_isect: torch.Tensor  # i32[capacity]
    global_from_compact_gid: torch.Tensor  # i32[N]
    tile_id_from_isect: Optional[torch.Tensor] = None  # i32[capacity], sorted (not in the reference's RenderAux)

    def tensors(self):
        return (self.projected_splats, self.uniforms_buffer, self.num_intersections, self.num_visible,
                self.final_index, self.cum_tiles_hit, self.tile_bins, self.compact_gid_from_isect,
                self.global_from_compact_gid, self.tile_id_from_isect)

    def read_num_visible(self) -> int:  # lib.rs:42-47
        return int(self.num_visible.item())

    def read_num_intersections(self) -> int:  # lib.rs:49-54
        return int(self.num_intersections.item())

    def read_tile_depth(self) -> torch.Tensor:  # lib.rs:56-62
        return self.tile_bins[..., 1] - self.tile_bins[..., 0]

    def _c(self) -> BrushAux:
        a = BrushAux()
        a.projected_splats = self.projected_splats.data_ptr()
        a.uniforms_buffer = self.uniforms_buffer.data_ptr()
        a.num_intersections = self.num_intersections.data_ptr()
        a.num_visible = self.num_visible.data_ptr()
        a.final_index = self.final_index.data_ptr()
        a.cum_tiles_hit = self.cum_tiles_hit.data_ptr()
        a.tile_bins = self.tile_bins.data_ptr()
        a.compact_gid_from_isect = self.compact_gid_from_isect.data_ptr()
        a.global_from_compact_gid = self.global_from_compact_gid.data_ptr()
        a.tile_id_from_isect = self.tile_id_from_isect.data_ptr() if self.tile_id_from_isect is not None else None
        a.isect_capacity = self.compact_gid_from_isect.numel()
        return a


_contexts = {}


def get_context(device) -> Context:
    """One library context per CUDA device (the analogue of the per-device compute client)."""
    idx = torch.device(device).index
    if idx is None:
        idx = torch.cuda.current_device()
    if idx not in _contexts:
        _contexts[idx] = Context(idx)
    return _contexts[idx]


def _check_dims(means, log_scales, quats, sh_coeffs, raw_opacity):
    """DimCheck, render.rs:74-79 (assert_eq! panics there, ValueError here)."""
    D = means.shape[0]
    ok = (means.dim() == 2 and means.shape[1] == 3 and tuple(log_scales.shape) == (D, 3)
          and tuple(quats.shape) == (D, 4) and sh_coeffs.dim() == 3 and sh_coeffs.shape[0] == D
          and sh_coeffs.shape[2] == 3 and tuple(raw_opacity.shape) == (D,))
    if not ok:
        raise ValueError("render_splats: dimension mismatch: means[D,3] log_scales[D,3] quats[D,4] "
                         f"sh_coeffs[D,C,3] raw_opacity[D]; got {tuple(means.shape)} {tuple(log_scales.shape)} "
                         f"{tuple(quats.shape)} {tuple(sh_coeffs.shape)} {tuple(raw_opacity.shape)}")
    for t in (means, log_scales, quats, sh_coeffs, raw_opacity):
        if t.dtype != torch.float32 or not t.is_cuda:
            raise ValueError("render_splats: inputs must be float32 CUDA tensors (no CPU fallback exists)")
    sh_degree_from_coeffs(sh_coeffs.shape[1])


def _stream_ptr(device) -> int:
    return torch.cuda.current_stream(device).cuda_stream


def render_forward(uniforms: BrushUniforms, means, log_scales, quats, sh_coeffs, raw_opacity, raster_u32: bool,
                   isect_capacity: Optional[int] = None, keep_tile_ids: bool = False, ctx: Optional[Context] = None,
                   reuse=None):
    """render.rs:55-323 through brush_render_forward.  Returns (out_img, RenderAux).  Async on the current stream.
    `reuse` = an (out_img, RenderAux) pair from an earlier call with the same shapes: its buffers are overwritten
    instead of allocating new ones (steady-state training loops)."""
    _check_dims(means, log_scales, quats, sh_coeffs, raw_opacity)
    dev = means.device
    ctx = ctx or get_context(dev)
    means, log_scales, quats = means.contiguous(), log_scales.contiguous(), quats.contiguous()
    sh_coeffs, raw_opacity = sh_coeffs.contiguous(), raw_opacity.contiguous()
    N, K = means.shape[0], sh_coeffs.shape[1]
    W, H = uniforms.img_size[0], uniforms.img_size[1]
    tx, ty = (W + TILE_WIDTH - 1) // TILE_WIDTH, (H + TILE_WIDTH - 1) // TILE_WIDTH
    if isect_capacity is None:
        isect_capacity = min(N * tx * ty, REFERENCE_MAX_INTERSECTS)  # render.rs:204-206
    cap = max(int(isect_capacity), 1)
    i32 = dict(dtype=torch.int32, device=dev)
    if reuse is not None:
        out, aux = reuse
        if (aux.projected_splats.shape[0] != N or aux.final_index.shape != (H, W)
                or aux.compact_gid_from_isect.numel() != cap or out.shape[-1] != (1 if raster_u32 else 4)):
            raise ValueError("render_forward: `reuse` buffers do not match this call's shapes")
        if poison_enabled():
            poison_(out, *aux.tensors())
        caux = aux._c()
        with torch.cuda.device(dev):
            rc = _lib.lib().brush_render_forward(
                ctx.handle, C.byref(uniforms), N, K, means.data_ptr(), log_scales.data_ptr(), quats.data_ptr(),
                sh_coeffs.data_ptr(), raw_opacity.data_ptr(), int(raster_u32), out.data_ptr(), C.byref(caux),
                C.c_void_p(_stream_ptr(dev)))
        ctx.check(rc)
        return out, aux
    aux = RenderAux(
        projected_splats=torch.empty((N, 9), dtype=torch.float32, device=dev),
        uniforms_buffer=torch.empty(28, **i32),
        num_intersections=torch.empty(1, **i32),
        num_visible=torch.empty(1, **i32),
        final_index=torch.empty((H, W), **i32),
        cum_tiles_hit=torch.empty(N, **i32),
        tile_bins=torch.empty((ty, tx, 2), **i32),
        compact_gid_from_isect=torch.empty(cap, **i32),
        global_from_compact_gid=torch.empty(N, **i32),
        tile_id_from_isect=torch.empty(cap, **i32) if keep_tile_ids else None,
    )
    if raster_u32:
        out = torch.empty((H, W, 1), dtype=torch.int32, device=dev)  # packed RGBA8 (render.rs:269-280)
    else:
        out = torch.empty((H, W, 4), dtype=torch.float32, device=dev)
    if poison_enabled():
        poison_(out, *aux.tensors())
    caux = aux._c()
    with torch.cuda.device(dev):
        rc = _lib.lib().brush_render_forward(
            ctx.handle, C.byref(uniforms), N, K, means.data_ptr(), log_scales.data_ptr(), quats.data_ptr(),
            sh_coeffs.data_ptr(), raw_opacity.data_ptr(), int(raster_u32), out.data_ptr(), C.byref(caux),
            C.c_void_p(_stream_ptr(dev)))
    ctx.check(rc)
    return out, aux


def render_backward(aux: RenderAux, means, log_scales, quats, sh_coeffs_shape, raw_opacity, out_img, v_out_img,
                    ctx: Optional[Context] = None, out=None):
    """render.rs:468-625 through brush_render_backward.  Returns the six dense gradients in the reference's
    registration order: means, xy_dummy, log_scales, quats, sh_coeffs, raw_opacity."""
    dev = means.device
    ctx = ctx or get_context(dev)
    N, K = means.shape[0], sh_coeffs_shape[1]
    f32 = dict(dtype=torch.float32, device=dev)
    if out is not None:  # caller-owned (e.g. views into one flat all-reduce buffer)
        v_means, v_xy, v_scales, v_quats, v_sh, v_opac = out
    else:
        v_means = torch.empty((N, 3), **f32)
        v_xy = torch.empty((N, 2), **f32)
        v_scales = torch.empty((N, 3), **f32)
        v_quats = torch.empty((N, 4), **f32)
        v_sh = torch.empty((N, K, 3), **f32)
        v_opac = torch.empty((N,), **f32)
    v_out_img = v_out_img.contiguous()
    if poison_enabled():
        poison_(v_means, v_xy, v_scales, v_quats, v_sh, v_opac)
    caux = aux._c()
    with torch.cuda.device(dev):
        rc = _lib.lib().brush_render_backward(
            ctx.handle, N, K, means.data_ptr(), log_scales.data_ptr(), quats.data_ptr(), raw_opacity.data_ptr(),
            out_img.data_ptr(), v_out_img.data_ptr(), C.byref(caux), v_means.data_ptr(), v_xy.data_ptr(),
            v_scales.data_ptr(), v_quats.data_ptr(), v_sh.data_ptr(), v_opac.data_ptr(),
            C.c_void_p(_stream_ptr(dev)))
    ctx.check(rc)
    return v_means, v_xy, v_scales, v_quats, v_sh, v_opac


class _RenderSplats(torch.autograd.Function):
    """The autodiff node: state = {aux, out_img, sh_degree} + 4 checkpointed inputs (render.rs:436-446)."""

    @staticmethod
    def forward(fctx, means, xy_dummy, log_scales, quats, sh_coeffs, raw_opacity, uniforms, isect_capacity,
                keep_tile_ids, holder):
        means, log_scales, quats = means.contiguous(), log_scales.contiguous(), quats.contiguous()
        raw_opacity = raw_opacity.contiguous()
        out, aux = render_forward(uniforms, means, log_scales, quats, sh_coeffs, raw_opacity, False,
                                  isect_capacity, keep_tile_ids)
        holder.append(aux)
        fctx.aux = aux
        fctx.sh_shape = tuple(sh_coeffs.shape)
        fctx.save_for_backward(means, log_scales, quats, raw_opacity, out)
        fctx.mark_non_differentiable()
        return out

    @staticmethod
    def backward(fctx, v_out):
        means, log_scales, quats, raw_opacity, out = fctx.saved_tensors
        g = render_backward(fctx.aux, means, log_scales, quats, fctx.sh_shape, raw_opacity, out, v_out)
        return g[0], g[1], g[2], g[3], g[4], g[5], None, None, None, None


def render_splats(cam: Camera, img_size, means, xy_grad_dummy, log_scales, quats, sh_coeffs, raw_opacity,
                  render_u32_buffer: bool, *, isect_capacity: Optional[int] = None, keep_tile_ids: bool = False):
    """brush_render::Backend::render_splats (lib.rs:66-86).  `img_size` = (w, h).  `quats` must already be
    normalised (Splats.render does it, gaussian_splats.rs:174-175).  Differentiable w.r.t. the six tensors
    unless render_u32_buffer (render.rs:282-284)."""
    u = make_uniforms(cam.world_to_local(), cam.focal(img_size), img_size, cam.center(img_size))
    tensors = (means, xy_grad_dummy, log_scales, quats, sh_coeffs, raw_opacity)
    needs_grad = torch.is_grad_enabled() and any(t is not None and t.requires_grad for t in tensors)
    if render_u32_buffer or not needs_grad:
        return render_forward(u, means.detach(), log_scales.detach(), quats.detach(), sh_coeffs.detach(),
                              raw_opacity.detach(), render_u32_buffer, isect_capacity, keep_tile_ids)
    _check_dims(means, log_scales, quats, sh_coeffs, raw_opacity)
    if xy_grad_dummy is None:
        xy_grad_dummy = torch.zeros((means.shape[0], 2), dtype=torch.float32, device=means.device)
    holder = []
    out = _RenderSplats.apply(means, xy_grad_dummy, log_scales, quats, sh_coeffs, raw_opacity, u, isect_capacity,
                              keep_tile_ids, holder)
    return out, holder[0]


def inverse_sigmoid(x: float) -> float:  # gaussian_splats.rs:38-40
    return math.log(x / (1.0 - x))


class Splats:
    """gaussian_splats.rs:24-36: the five parameters + the xy dummy that carries screen-space gradients."""

    def __init__(self, means, sh_coeffs, rotation, raw_opacity, log_scales):  # from_data, :138-157
        def p(t):
            return t.detach().clone().float().contiguous().requires_grad_(True)

        self.means, self.sh_coeffs, self.rotation = p(means), p(sh_coeffs), p(rotation)
        self.raw_opacity, self.log_scales = p(raw_opacity), p(log_scales)
        self.xys_dummy = torch.zeros((self.means.shape[0], 2), dtype=torch.float32, device=self.means.device,
                                     requires_grad=True)

    @classmethod
    def from_views(cls, means, sh_coeffs, rotation, raw_opacity, log_scales) -> "Splats":
        """Wraps existing storage WITHOUT copying (the peer exchange keeps the parameters in a peer-mapped flat
        buffer that brush_dp_adam_step updates in place)."""
        self = cls.__new__(cls)

        def p(t):
            assert t.dtype == torch.float32 and t.is_contiguous()
            return t.detach().requires_grad_(True)

        self.means, self.sh_coeffs, self.rotation = p(means), p(sh_coeffs), p(rotation)
        self.raw_opacity, self.log_scales = p(raw_opacity), p(log_scales)
        self.xys_dummy = torch.zeros((self.means.shape[0], 2), dtype=torch.float32, device=self.means.device,
                                     requires_grad=True)
        return self

    @classmethod
    def from_safetensors(cls, tensors: dict, device) -> "Splats":  # gaussian_splats.rs:208-223
        def t(name):
            return torch.as_tensor(np.asarray(tensors[name]), dtype=torch.float32, device=device)

        return cls(t("means"), t("coeffs"), t("quats"), t("opacities"), t("scales"))

    def params(self):
        return [self.means, self.sh_coeffs, self.rotation, self.raw_opacity, self.log_scales]

    def render(self, camera: Camera, img_size, render_u32_buffer: bool, **kw):  # gaussian_splats.rs:167-188
        rot = self.rotation
        norm_rot = rot / torch.sqrt(torch.sum(rot * rot, dim=1, keepdim=True))
        return render_splats(camera, img_size, self.means, self.xys_dummy, self.log_scales, norm_rot,
                             self.sh_coeffs, self.raw_opacity, render_u32_buffer, **kw)

    def opacity(self):
        return torch.sigmoid(self.raw_opacity)

    def scales(self):
        return torch.exp(self.log_scales)

    def num_splats(self) -> int:
        return self.means.shape[0]

    def norm_rotations(self):  # gaussian_splats.rs:202-206
        with torch.no_grad():
            n = torch.clamp_min(torch.sqrt(torch.sum(self.rotation ** 2, dim=1, keepdim=True)), 1e-6)
            self.rotation.div_(n)
