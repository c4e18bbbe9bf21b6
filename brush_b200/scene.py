"""Synthetic random-Gaussian scenes for the BASELINE.json configs.  numpy, seeded,
float32.  Camera = the reference bench camera (benches/render_bench.rs:163-174): position (0,0,-8), identity
rotation, fov_x = pi/2, fov_y from the same focal."""
from __future__ import annotations

import math
from dataclasses import dataclass

import numpy as np

CONFIGS = {
    # name: (N, width, height, sh_degree, sigma_px range)
    "cfg1_16k_256": (16_384, 256, 256, 0, (0.5, 4.0)),
    "cfg2_100k_1024": (100_000, 1024, 1024, 0, (0.5, 4.0)),
    "cfg3_300k_800": (300_000, 800, 800, 3, (0.5, 4.0)),
    "cfg4_1m_1080p": (1_000_000, 1920, 1080, 3, (0.5, 4.0)),
    "cfg5_6m_4k": (6_000_000, 3840, 2160, 3, (2.0, 12.0)),
}


@dataclass
class Scene:
    means: np.ndarray
    log_scales: np.ndarray
    quats: np.ndarray
    sh_coeffs: np.ndarray
    raw_opacity: np.ndarray
    width: int
    height: int
    sh_degree: int


def bench_camera_params(width: int, height: int):
    """(position, rotation xyzw, fov_x, fov_y) of the reference bench camera."""
    fov = math.pi * 0.5
    focal = 0.5 * width / math.tan(fov * 0.5)
    fov_x = 2.0 * math.atan(width / (2.0 * focal))
    fov_y = 2.0 * math.atan(height / (2.0 * focal))
    return (0.0, 0.0, -8.0), (0.0, 0.0, 0.0, 1.0), fov_x, fov_y


def ring_camera_params(width: int, height: int, k: int, n_views: int, radius: float = 8.0):
    """Camera k of n on a ring of radius 8 around the origin, looking at it (multi-view runs)."""
    ang = 2.0 * math.pi * k / max(n_views, 1)
    pos = (-radius * math.sin(ang), 0.0, -radius * math.cos(ang))
    half = 0.5 * ang  # rotation about +y by ang
    rot = (0.0, math.sin(half), 0.0, math.cos(half))
    _, _, fov_x, fov_y = bench_camera_params(width, height)
    return pos, rot, fov_x, fov_y


def make_scene(n: int, width: int, height: int, sh_degree: int, seed: int, sigma_px=(0.5, 4.0)) -> Scene:
    rng = np.random.default_rng(seed)
    focal = 0.5 * width  # fov_x = 90 deg
    tan_x, tan_y = 1.0, height / width
    z_v = rng.uniform(2.0, 14.0, n)
    x = rng.uniform(-1.1, 1.1, n) * z_v * tan_x
    y = rng.uniform(-1.1, 1.1, n) * z_v * tan_y
    means = np.stack([x, y, z_v - 8.0], 1)  # camera at z=-8 looking down +z
    s_px = np.exp(rng.uniform(math.log(sigma_px[0]), math.log(sigma_px[1]), n))
    aniso = rng.uniform(0.5, 2.0, (n, 3))
    scales = (s_px * z_v / focal)[:, None] * aniso
    u, v, w = rng.uniform(0, 1, n), rng.uniform(0, 2 * math.pi, n), rng.uniform(0, 2 * math.pi, n)
    quats = np.stack([np.sqrt(1 - u) * np.sin(v), np.sqrt(1 - u) * np.cos(v), np.sqrt(u) * np.sin(w),
                      np.sqrt(u) * np.cos(w)], 1)
    p = rng.uniform(0.05, 0.95, n)
    raw_opac = np.log(p / (1 - p))
    k = (sh_degree + 1) ** 2
    sh = rng.uniform(-0.25, 0.25, (n, k, 3))
    sh[:, 0, :] = rng.uniform(-1.0, 1.0, (n, 3))
    f = np.float32
    return Scene(means.astype(f), np.log(scales).astype(f), quats.astype(f), sh.astype(f), raw_opac.astype(f),
                 width, height, sh_degree)


def make_config_scene(name: str, n_override: int | None = None) -> Scene:
    n, w, h, deg, sig = CONFIGS[name]
    idx = list(CONFIGS).index(name) + 1
    return make_scene(n_override or n, w, h, deg, 1000 + idx, sig)
