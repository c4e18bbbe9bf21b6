"""The C ABI driven from C, the way the Rust FFI crate drives it (rust/brush-b200-sys): tests/abi_c/test_abi.c is
compiled with gcc, linked against libbrush_b200.so + the CUDA runtime only, and run on the reference's golden cases
with raw cudaMalloc buffers.  Its outputs are compared here at the reference's own tolerances (render.rs:815-830).
The compile-and-link half runs without a GPU."""
import math
import os
import subprocess

import numpy as np
import pytest

from helpers import all_close, crab, load_case, normalize_quats, quat_norm_vjp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SRC = os.path.join(ROOT, "tests", "abi_c", "test_abi.c")
CUDA = os.environ.get("CUDA_HOME", "/usr/local/cuda")


def _compile(tmp_path):
    from brush_b200 import build

    so = build.build()
    exe = str(tmp_path / "test_abi")
    libdir = os.path.dirname(so)
    cmd = ["gcc", "-std=c11", "-O1", "-Wall", "-Werror", SRC, "-I", os.path.join(ROOT, "include"),
           "-I", os.path.join(CUDA, "include"), "-L", libdir, "-lbrush_b200", "-L", os.path.join(CUDA, "lib64"), "-lcudart",
           "-lm", f"-Wl,-rpath,{libdir}", f"-Wl,-rpath,{os.path.join(CUDA, 'lib64')}", "-o", exe]
    r = subprocess.run(cmd, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    return exe


def test_c_caller_compiles_and_links_against_the_header(tmp_path):
    """The header is valid C11 (not only C++), every entry point the C caller uses resolves at link time, and the
    binary's only product dependency is libbrush_b200.so."""
    exe = _compile(tmp_path)
    needed = subprocess.run(["readelf", "-d", exe], capture_output=True, text=True).stdout
    assert "libbrush_b200.so" in needed and "libcudart" in needed and "torch" not in needed


@pytest.mark.gpu
@pytest.mark.parametrize("name", ["tiny_case", "basic_case"])
def test_c_caller_reproduces_the_reference_golden_case(tmp_path, name):
    exe = _compile(tmp_path)
    d = load_case(name)
    h, w, _ = d["out_img"].shape
    n, k = d["means"].shape[0], d["coeffs"].shape[1]
    qhat, qn = normalize_quats(d["quats"].astype(np.float32))
    work = tmp_path / name
    work.mkdir()
    for fname, arr in (("means", d["means"]), ("log_scales", d["scales"]), ("quats", qhat), ("sh_coeffs", d["coeffs"]),
                       ("raw_opacity", d["opacities"]), ("target", crab())):
        np.ascontiguousarray(arr, dtype=np.float32).tofile(str(work / (fname + ".f32")))
    r = subprocess.run([exe, str(work), str(n), str(k), str(w), str(h)], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout + r.stderr
    rd = lambda f, shape, dt=np.float32: np.fromfile(str(work / (f + ".bin")), dtype=dt).reshape(shape)
    counts = open(work / "counts.txt").read().split()
    nv, ni = int(counts[0]), int(counts[1])
    assert 0 < nv <= n and ni > 0 and int(counts[2]) > 10
    perm = rd("out_gfc", (n,), np.int32)[:nv]
    proj = rd("out_projected", (n, 9))
    assert all_close(proj[:nv, 0:2], d["xys"][perm], 1e-4, 1e-10)
    assert all_close(proj[:nv, 2:5], d["conics"][perm], 1e-4, 5e-7)
    img = rd("out_img", (h, w, 4))
    assert all_close(img[..., :3], d["out_img"], 1e-4, 1e-9)
    assert np.array_equal(rd("out_img_host", (h, w, 4)), img)  # brush_render_forward_host: same kernels, same bits
    # the C program forms d loss / d out of loss = mean((out_rgb - crab)^2) from its own render (render.rs:786-789)
    assert all_close(rd("out_v_xy", (n, 2)), d["v_xy"], 1e-4, 1e-9)
    assert all_close(rd("out_v_opac", (n,)), d["v_opacities"], 1e-4, 1e-10)
    assert all_close(rd("out_v_sh", (n, k, 3)), d["v_coeffs"], 1e-4, 1e-9)
    assert all_close(rd("out_v_log_scales", (n, 3)), d["v_scales"], 1e-4, 1e-9)
    assert all_close(rd("out_v_means", (n, 3)), d["v_means"], 1e-4, 1e-9)
    # v_quats is w.r.t. the NORMALISED rotation here; the golden value is w.r.t. the raw one (loose in the reference too)
    assert all_close(quat_norm_vjp(qhat, qn, rd("out_v_quats", (n, 4))), d["v_quats"], 1e-1, 1e-1)
    stage_ms = [float(x) for x in counts[3:]]
    assert len(stage_ms) == 10 and all(x >= 0 for x in stage_ms) and stage_ms[7] > 0  # Rasterize was timed
