"""The callers and data formats either side of the path: the view feed, the NeRF-synthetic
camera convention and the evaluation loop.

    image_to_tensor            == crates/brush-train/src/image.rs:8-18
    SceneView / Scene.bounds   == crates/brush-train/src/scene.rs:13-53, brush-render/src/bounding_box.rs:8-13
    camera_from_nerf_transform == crates/brush-dataset/src/formats/nerf_synthetic.rs:55-68,86-92
    read_nerf_synthetic        == nerf_synthetic.rs:26-140 (directory instead of zip archive)
    SceneLoader                == crates/brush-dataset/src/scene_loader.rs:11-67 (5 batches prefetched)
    eval_stats                 == crates/brush-train/src/eval.rs:27-77 (PSNR in RGB, SSIM window 11)

Host-side logic (numpy + threads); images reach the device as pinned-host tensors that SplatTrainer.step copies on
its stream.  With world > 1 every rank builds the same loader (same seed) and takes its own slot of each draw."""
from __future__ import annotations

import json
import math
import os
import queue
import threading
from dataclasses import dataclass
from typing import List, Optional

import numpy as np
import torch

from .render import Camera, Splats, focal_to_fov, fov_to_focal
from .train import SceneBatch


def image_to_tensor(image: np.ndarray) -> torch.Tensor:
    """image.rs:8-18: u8 [H,W,3|4] -> f32 [H,W,3|4] in [0,1] (image crate's to_rgb32f: v / 255)."""
    a = np.asarray(image)
    if a.dtype == np.uint8:
        a = a.astype(np.float32) / np.float32(255.0)
    if a.ndim != 3 or a.shape[2] not in (3, 4):
        raise ValueError("image must be [H, W, 3] or [H, W, 4]")
    return torch.from_numpy(np.ascontiguousarray(a, dtype=np.float32))


@dataclass
class SceneView:  # scene.rs:13-18
    name: str
    camera: Camera
    image: np.ndarray  # u8 [H, W, 3|4]


def _rotate(q_xyzw: np.ndarray, v: np.ndarray) -> np.ndarray:
    x, y, z, w = q_xyzw
    u = np.array([x, y, z])
    return v + 2.0 * np.cross(u, np.cross(u, v) + w * v)


class Scene:  # scene.rs:20-53
    def __init__(self, views: List[SceneView]):
        self.views = list(views)

    def bounds(self, cam_near: float, cam_far: float):
        """(center, extent) of the cameras' positions pushed `near` / `far` along their view axis (scene.rs:42-53;
        extent = half the box size, bounding_box.rs:8-13)."""
        lo, hi = np.full(3, np.inf), np.full(3, -np.inf)
        for v in self.views:
            fwd = _rotate(v.camera.rotation / np.linalg.norm(v.camera.rotation), np.array([0.0, 0.0, 1.0]))
            for d in (cam_near, cam_far):
                p = v.camera.position + fwd * d
                lo, hi = np.minimum(lo, p), np.maximum(hi, p)
        return (hi + lo) / 2.0, (hi - lo) / 2.0

    def scene_extent(self) -> float:  # scene_loader.rs:21
        return float(self.bounds(0.0, 0.0)[1].max())


def _quat_from_rotmat(m: np.ndarray) -> np.ndarray:
    """glam Quat::from_mat3 (x, y, z, w) of a pure rotation."""
    t = np.trace(m)
    if t > 0:
        s = math.sqrt(t + 1.0) * 2
        return np.array([(m[2, 1] - m[1, 2]) / s, (m[0, 2] - m[2, 0]) / s, (m[1, 0] - m[0, 1]) / s, 0.25 * s])
    i = int(np.argmax(np.diag(m)))
    j, k = (i + 1) % 3, (i + 2) % 3
    s = math.sqrt(1.0 + m[i, i] - m[j, j] - m[k, k]) * 2
    q = np.zeros(4)
    q[i] = 0.25 * s
    q[j] = (m[j, i] + m[i, j]) / s
    q[k] = (m[k, i] + m[i, k]) / s
    q[3] = (m[k, j] - m[j, k]) / s
    return q


def camera_from_nerf_transform(transform_matrix, fov_x: float, width: int, height: int) -> Camera:
    """nerf_synthetic.rs:55-68,86-92.  `transform_matrix` = the 4x4 camera-to-world rows of transforms_*.json.
    Flip the camera's y and z axes (OpenGL -> y-down, z-forward), then rotate the world +90 degrees about x."""
    m = np.asarray(transform_matrix, dtype=np.float64).reshape(4, 4).copy()
    m[:, 1] *= -1.0  # transform.y_axis *= -1 (columns of the camera-to-world matrix)
    m[:, 2] *= -1.0
    c, s = math.cos(math.pi / 2), math.sin(math.pi / 2)
    rx = np.array([[1, 0, 0, 0], [0, c, -s, 0], [0, s, c, 0], [0, 0, 0, 1]], dtype=np.float64)
    m = rx @ m
    rot = m[:3, :3]
    scale = np.linalg.norm(rot, axis=0)  # to_scale_rotation_translation
    q = _quat_from_rotmat(rot / scale)
    fov_y = focal_to_fov(fov_to_focal(fov_x, width), height)  # :86
    return Camera(m[:3, 3], q, fov_x, fov_y, (0.5, 0.5))


def read_nerf_synthetic(root: str, split: str = "train", max_frames: Optional[int] = None) -> Scene:
    """transforms_{split}.json + the PNGs it names -> Scene (nerf_synthetic.rs:26-99)."""
    from PIL import Image

    with open(os.path.join(root, f"transforms_{split}.json")) as f:
        meta = json.load(f)
    fov_x = float(meta["camera_angle_x"])
    views = []
    for frame in meta["frames"][:max_frames]:
        path = os.path.join(root, frame["file_path"] + ".png")
        img = np.asarray(Image.open(path))
        if img.ndim == 2:
            img = np.repeat(img[..., None], 3, 2)
        cam = camera_from_nerf_transform(frame["transform_matrix"], fov_x, img.shape[1], img.shape[0])
        views.append(SceneView(path, cam, np.ascontiguousarray(img)))
    return Scene(views)


class SceneLoader:
    """scene_loader.rs:11-67: a background producer keeps up to 5 batches ready; each batch = `batch_size` views
    drawn uniformly with replacement from a seeded generator, images stacked [B, H, W, C] (pinned host memory here,
    so the H2D copy can run on the training stream).  rank / world: every rank draws the same global batch of
    world * batch_size views and keeps its own slice, so one step covers distinct views on distinct GPUs."""

    PREFETCH = 5

    def __init__(self, scene: Scene, batch_size: int, seed: int, rank: int = 0, world: int = 1, pin: bool = True,
                 quantized: bool = False):
        if not scene.views:
            raise ValueError("SceneLoader needs at least one view")
        self.scene, self.batch_size = scene, batch_size
        self._extent = scene.scene_extent()
        self._rng = np.random.default_rng(seed)
        self._rank, self._world = rank, world
        self._pin = pin and torch.cuda.is_available()
        # quantized: batches keep the 8-bit pixels; SplatTrainer uploads the bytes and applies image_to_tensor's
        # /255 on the device (a quarter of the PCIe traffic per step)
        self._quantized = quantized
        self._q: "queue.Queue[SceneBatch]" = queue.Queue(maxsize=self.PREFETCH)
        self._stop = threading.Event()
        self._thread = threading.Thread(target=self._produce, daemon=True)
        self._thread.start()

    def _make_batch(self) -> SceneBatch:
        idx = self._rng.integers(0, len(self.scene.views), self.batch_size * self._world)
        mine = idx[self._rank * self.batch_size:(self._rank + 1) * self.batch_size]
        views = [self.scene.views[i] for i in mine]
        if self._quantized and all(np.asarray(v.image).dtype == np.uint8 for v in views):
            imgs = torch.stack([torch.from_numpy(np.ascontiguousarray(v.image)) for v in views], 0)
        else:
            imgs = torch.stack([image_to_tensor(v.image) for v in views], 0)
        if self._pin:
            imgs = imgs.pin_memory()
        return SceneBatch(imgs, [v.camera for v in views], self._extent)

    def _produce(self):
        while not self._stop.is_set():
            batch = self._make_batch()
            while not self._stop.is_set():
                try:
                    self._q.put(batch, timeout=0.1)
                    break
                except queue.Full:
                    continue

    def next_batch(self) -> SceneBatch:
        return self._q.get()

    def close(self):
        self._stop.set()


@dataclass
class EvalView:  # eval.rs:11-20
    view: SceneView
    rendered: torch.Tensor
    psnr: float
    ssim: float
    aux: object


def eval_stats(splats: Splats, eval_scene: Scene, num_frames: Optional[int], rng: np.random.Generator, device):
    """eval.rs:27-77.  PSNR = 10 log10(1 / mse) over RGB; SSIM = the train-step statistic (window 11), computed by
    the loss kernel with ssim_weight = 1 (its loss value is -ssim)."""
    import ctypes as C

    from . import _lib
    from .render import _stream_ptr, get_context

    n = len(eval_scene.views)
    idx = list(range(n)) if num_frames is None else sorted(rng.choice(n, min(num_frames, n), replace=False).tolist())
    ctx, L = get_context(device), _lib.lib()
    out = []
    for i in idx:
        view = eval_scene.views[i]
        gt = image_to_tensor(view.image[..., :3]).to(device)  # to_rgb8, eval.rs:49
        h, w = gt.shape[:2]
        with torch.no_grad():
            rendered, aux = splats.render(view.camera, (w, h), False)
        rgb = rendered[..., :3]
        mse = torch.mean((rgb - gt) ** 2)
        psnr = float(10.0 * torch.log10(1.0 / mse))
        v_out, loss = torch.empty((h, w, 4), device=device), torch.zeros(1, device=device)
        with torch.cuda.device(device):
            ctx.check(L.brush_image_loss_backward(ctx.handle, rendered.contiguous().data_ptr(), gt.data_ptr(), h, w, 3,
                                                  1.0, 11, v_out.data_ptr(), loss.data_ptr(),
                                                  C.c_void_p(_stream_ptr(torch.device(device)))))
        out.append(EvalView(view, rgb, psnr, -float(loss), aux))
    return out
