"""Shared test helpers: golden-fixture loading and the reference test camera."""
import math
import os

import numpy as np

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def load_case(name):
    from safetensors.numpy import load_file

    return load_file(os.path.join(GOLDEN, name + ".safetensors"))


def crab():
    return np.load(os.path.join(GOLDEN, "crab_rgb8.npy")).astype(np.float32) / 255.0


def reference_camera(w, h):
    """Camera of test_reference (render.rs:734-746): pos (0,0,-8), identity, fov 90 deg on width.

    Returns (viewmat row-major 4x4 f32, (fx, fy), (cx, cy))."""
    fov = math.pi * 0.5
    focal = 0.5 * w / math.tan(fov * 0.5)            # camera.rs:52-54
    fov_x = 2.0 * math.atan(w / (2.0 * focal))       # camera.rs:57-59
    fov_y = 2.0 * math.atan(h / (2.0 * focal))
    fx = 0.5 * w / math.tan(fov_x * 0.5)
    fy = 0.5 * h / math.tan(fov_y * 0.5)
    vm = np.eye(4, dtype=np.float32)
    vm[2, 3] = 8.0  # inverse of translation (0,0,-8)
    return vm, (fx, fy), (0.5 * w, 0.5 * h)


def all_close(a, b, rtol, atol):
    """Burn's all_close: |a-b| <= atol + rtol*|b| elementwise."""
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    return bool((np.abs(a - b) <= atol + rtol * np.abs(b)).all())


def normalize_quats(q):
    n = np.linalg.norm(q, axis=1, keepdims=True).astype(np.float32)
    return (q / n).astype(np.float32), n


def quat_norm_vjp(qhat, n, v):
    """VJP of q -> q/|q| (gaussian_splats.rs:174-175, done by Burn autodiff in the reference)."""
    return (v - qhat * (qhat * v).sum(1, keepdims=True)) / n
