"""TEST INFRASTRUCTURE ONLY -- ctypes binding of the CPU oracle (oracle/libbrush_oracle.so).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / ``--impl reference`` legs
may import this module.  The product package (brush_b200/) never does.

The oracle restates the reference shaders (see brush_oracle.cpp for file:line citations);
this wrapper mirrors the host sequencing of crates/brush-render/src/render.rs with numpy
arrays in place of Burn tensors.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
from dataclasses import dataclass

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "libbrush_oracle.so")
TILE_WIDTH = 16


class Uniforms(C.Structure):
    """helpers.wgsl:7-30 (112 bytes)."""

    _fields_ = [
        ("viewmat", C.c_float * 16),
        ("focal", C.c_float * 2),
        ("img_size", C.c_uint32 * 2),
        ("tile_bounds", C.c_uint32 * 2),
        ("pixel_center", C.c_float * 2),
        ("sh_degree", C.c_uint32),
        ("num_visible", C.c_uint32),
        ("total_splats", C.c_uint32),
        ("padding", C.c_uint32),
    ]


assert C.sizeof(Uniforms) == 112


def build(force: bool = False) -> str:
    src = [os.path.join(_HERE, f) for f in ("brush_oracle.cpp", "oracle_math.h", "Makefile")]
    if force or not os.path.exists(_SO) or any(os.path.getmtime(s) > os.path.getmtime(_SO) for s in src):
        subprocess.check_call(["make", "-C", _HERE, "-s"])
    return _SO


_TIMED_DIR = os.path.join(_HERE, "_timed")
_TIMED_SO = os.path.join(_TIMED_DIR, "libbrush_oracle_timed.so")


def _cpu_tag() -> str:
    """Identifies the CPU a -march=native build is valid for (model name + ISA flags of /proc/cpuinfo)."""
    import hashlib

    model, flags = "", ""
    try:
        with open("/proc/cpuinfo") as f:
            for line in f:
                if line.startswith("model name") and not model:
                    model = line.split(":", 1)[1].strip()
                elif line.startswith("flags") and not flags:
                    flags = " ".join(sorted(line.split(":", 1)[1].split()))
                if model and flags:
                    break
    except OSError:
        pass
    return hashlib.sha256((model + "|" + flags).encode()).hexdigest()[:16]


def use_timed() -> str:
    """Switches this module to the TIMED build (oracle/Makefile `timed`: -O3 -march=native, thread-private f32
    accumulators, 8-bit sort digits) -- the CPU baseline bench.py times.  Built on the machine that runs it: a copy
    built for another CPU (the tag differs) is rebuilt, never executed."""
    global _lib
    tag_path = os.path.join(_TIMED_DIR, "cpu.tag")
    tag = _cpu_tag()
    src = [os.path.join(_HERE, f) for f in ("brush_oracle.cpp", "oracle_math.h", "Makefile")]
    stale = (not os.path.exists(_TIMED_SO) or not os.path.exists(tag_path) or open(tag_path).read().strip() != tag
             or any(os.path.getmtime(s_) > os.path.getmtime(_TIMED_SO) for s_ in src))
    if stale:
        if os.path.exists(_TIMED_SO):
            os.remove(_TIMED_SO)
        subprocess.check_call(["make", "-C", _HERE, "-s", "timed"])
        with open(tag_path, "w") as f:
            f.write(tag + "\n")
    _lib = C.CDLL(_TIMED_SO)
    _bind(_lib)
    return _TIMED_SO


def _bind(l):
    l.oracle_det_expf.restype = C.c_float
    l.oracle_det_expf.argtypes = [C.c_float]
    l.oracle_det_logf.restype = C.c_float
    l.oracle_det_logf.argtypes = [C.c_float]


_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(_SO):
            build()
        _lib = C.CDLL(_SO)
        _bind(_lib)
    return _lib


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def _f32(a):
    return np.ascontiguousarray(a, dtype=np.float32)


def _u32(a):
    return np.ascontiguousarray(a, dtype=np.uint32)


def set_threads(n: int):
    lib().oracle_set_threads(int(n))


def max_threads() -> int:
    return int(lib().oracle_max_threads())


STAGE_NAMES = ("ProjectSplats", "DepthSort", "ProjectVisibile", "PrefixSum", "MapGaussiansToIntersect", "Tile sort",
               "GetTileBinEdges", "Rasterize", "RasterizeBackwards", "GatherGrads", "ProjectBackwards")


def stage_seconds() -> dict:
    """Wall-clock seconds per stage of the last render_forward / render_backward (the reference's span names)."""
    arr = (C.c_double * 11)()
    lib().oracle_stage_seconds(arr)
    return dict(zip(STAGE_NAMES, [float(x) for x in arr]))


def make_uniforms(viewmat, focal, img_size, pixel_center, sh_degree=0, total_splats=0) -> Uniforms:
    """viewmat: 4x4 world->camera matrix (row-major numpy, i.e. viewmat[r, c]); stored column-major."""
    u = Uniforms()
    vm = np.asarray(viewmat, dtype=np.float32)
    cols = vm.T.reshape(-1)  # column-major flatten
    for i in range(16):
        u.viewmat[i] = float(cols[i])
    u.focal[0], u.focal[1] = float(np.float32(focal[0])), float(np.float32(focal[1]))
    u.img_size[0], u.img_size[1] = int(img_size[0]), int(img_size[1])
    u.tile_bounds[0] = (int(img_size[0]) + TILE_WIDTH - 1) // TILE_WIDTH
    u.tile_bounds[1] = (int(img_size[1]) + TILE_WIDTH - 1) // TILE_WIDTH
    u.pixel_center[0], u.pixel_center[1] = float(np.float32(pixel_center[0])), float(np.float32(pixel_center[1]))
    u.sh_degree = sh_degree
    u.total_splats = total_splats
    return u


def radix_argsort_u32(keys, vals, n=None, bits=32):
    keys, vals = _u32(keys), _u32(vals)
    n = len(keys) if n is None else n
    ko, vo = np.zeros_like(keys), np.zeros_like(vals)
    rc = lib().oracle_radix_argsort_u32(_p(keys), _p(vals), C.c_uint32(n), C.c_uint32(bits), _p(ko), _p(vo))
    assert rc == 0
    return ko, vo


def radix_argsort_u64(keys, vals, n=None, bits=64):
    keys = np.ascontiguousarray(keys, dtype=np.uint64)
    vals = _u32(vals)
    n = len(keys) if n is None else n
    ko, vo = np.zeros_like(keys), np.zeros_like(vals)
    rc = lib().oracle_radix_argsort_u64(_p(keys), _p(vals), C.c_uint32(n), C.c_uint32(bits), _p(ko), _p(vo))
    assert rc == 0
    return ko, vo


def prefix_sum_u32(x):
    x = _u32(x)
    out = np.zeros_like(x)
    lib().oracle_prefix_sum_u32(_p(x), C.c_uint32(len(x)), _p(out))
    return out


@dataclass
class Aux:
    """RenderAux (crates/brush-render/src/lib.rs:20-33) as numpy arrays."""

    projected_splats: np.ndarray
    uniforms: Uniforms
    num_intersections: int
    num_visible: int
    final_index: np.ndarray
    cum_tiles_hit: np.ndarray
    tile_bins: np.ndarray
    compact_gid_from_isect: np.ndarray
    global_from_compact_gid: np.ndarray
    tile_id_from_isect: np.ndarray
    status: int = 0


def render_forward(u: Uniforms, means, log_scales, quats, coeffs, raw_opac, raster_u32=False, capacity=None):
    """render.rs:55-323.  Returns (out_img, Aux)."""
    means, log_scales, quats = _f32(means), _f32(log_scales), _f32(quats)
    coeffs, raw_opac = _f32(coeffs), _f32(raw_opac)
    N, K = means.shape[0], coeffs.shape[1]
    assert means.shape == (N, 3) and log_scales.shape == (N, 3) and quats.shape == (N, 4)
    assert coeffs.shape == (N, K, 3) and raw_opac.shape == (N,)
    W, H = u.img_size[0], u.img_size[1]
    tx, ty = (W + 15) // 16, (H + 15) // 16
    if capacity is None:
        capacity = min(N * tx * ty, 128 * 65535)  # render.rs:204-206
    capacity = max(int(capacity), 1)
    out = np.zeros((H, W), np.uint32) if raster_u32 else np.zeros((H, W, 4), np.float32)
    projected = np.full((N, 9), -12345.0, np.float32)
    final_index = np.zeros((H, W), np.uint32)
    cum = np.zeros(N, np.uint32)
    bins = np.zeros((ty, tx, 2), np.uint32)
    cg = np.zeros(capacity, np.uint32)
    tid = np.zeros(capacity, np.uint32)
    gfc = np.zeros(N, np.uint32)
    uo = Uniforms()
    ni, nv = C.c_uint32(0), C.c_uint32(0)
    rc = lib().oracle_render_forward(
        C.byref(u), C.c_uint32(N), C.c_uint32(K), _p(means), _p(log_scales), _p(quats), _p(coeffs), _p(raw_opac),
        C.c_int(int(raster_u32)), _p(out), _p(projected), C.byref(uo), C.byref(ni), C.byref(nv), _p(final_index),
        _p(cum), _p(bins), _p(cg), _p(gfc), _p(tid), C.c_uint64(capacity))
    if rc == -1:
        raise ValueError(f"Invalid nr. of sh bases {K}")
    return out, Aux(projected, uo, ni.value, nv.value, final_index, cum, bins, cg, gfc, tid, rc)


def render_backward(aux: Aux, means, log_scales, quats, raw_opac, K, out_img, v_out_img):
    """render.rs:468-625.  Returns dict of dense gradients."""
    means, log_scales, quats, raw_opac = _f32(means), _f32(log_scales), _f32(quats), _f32(raw_opac)
    out_img, v_out_img = _f32(out_img), _f32(v_out_img)
    N = means.shape[0]
    g = dict(v_means=np.zeros((N, 3), np.float32), v_xy=np.zeros((N, 2), np.float32),
             v_log_scales=np.zeros((N, 3), np.float32), v_quats=np.zeros((N, 4), np.float32),
             v_coeffs=np.zeros((N, K, 3), np.float32), v_raw_opac=np.zeros(N, np.float32))
    rc = lib().oracle_render_backward(
        C.byref(aux.uniforms), C.c_uint32(N), C.c_uint32(K), _p(means), _p(log_scales), _p(quats), _p(raw_opac),
        _p(out_img), _p(v_out_img), _p(aux.projected_splats), _p(aux.final_index), _p(aux.tile_bins),
        _p(aux.compact_gid_from_isect), _p(aux.global_from_compact_gid), _p(g["v_means"]), _p(g["v_xy"]),
        _p(g["v_log_scales"]), _p(g["v_quats"]), _p(g["v_coeffs"]), _p(g["v_raw_opac"]))
    assert rc == 0
    return g
