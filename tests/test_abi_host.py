"""CPU-only checks: the C-ABI library loads and exports every symbol include/brush_b200.h declares, and the
host-side mirror (Camera, DimCheck, sh-degree validation) behaves like the reference.  No compute calls."""
import math
import os
import re

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def built():
    from brush_b200 import build

    return build.build()


def test_header_symbols_exported(built):
    import ctypes

    hdr = open(os.path.join(ROOT, "include", "brush_b200.h")).read()
    names = set(re.findall(r"\b(brush_[a-z0-9_]+)\s*\(", hdr))
    names.discard("brush_ctx")
    assert {"brush_render_forward", "brush_render_backward", "brush_radix_argsort_u32", "brush_radix_argsort_u64",
            "brush_prefix_sum_u32", "brush_ctx_create"} <= names
    lib = ctypes.CDLL(built)
    for n in names:
        assert hasattr(lib, n), f"{n} declared in the header but not exported"
    from brush_b200 import _lib

    assert set(_lib.exported_symbols()) <= names
    assert _lib.lib().brush_version().decode().startswith("brush_b200")


def test_uniform_layout():
    import ctypes

    from brush_b200._lib import BrushUniforms

    assert ctypes.sizeof(BrushUniforms) == 112
    assert BrushUniforms.num_visible.offset == 25 * 4  # render.rs:145-149


def test_camera_matches_reference_test_camera():
    from brush_b200 import Camera, fov_to_focal, focal_to_fov

    w, h = 123, 82
    focal = fov_to_focal(math.pi * 0.5, w)
    cam = Camera((0, 0, -8), (0, 0, 0, 1), focal_to_fov(focal, w), focal_to_fov(focal, h))
    vm = cam.world_to_local()
    expect = np.eye(4, dtype=np.float32)
    expect[2, 3] = 8.0
    assert np.array_equal(vm, expect)
    fx, fy = cam.focal((w, h))
    assert abs(fx - 61.5) < 1e-4 and abs(fy - 61.5) < 1e-4
    assert cam.center((w, h)) == (61.5, 41.0)
    # rotated camera: world_to_local is the rigid inverse
    cam2 = Camera((1, 2, 3), (0.0, math.sin(0.3), 0.0, math.cos(0.3)), 1.0, 1.0)
    assert np.allclose(cam2.world_to_local().astype(np.float64) @ cam2.local_to_world(), np.eye(4), atol=1e-6)


def test_dim_check_and_no_cpu_fallback():
    from brush_b200 import Camera, render_splats, sh_degree_from_coeffs

    with pytest.raises(ValueError, match="Invalid nr. of sh bases"):
        sh_degree_from_coeffs(5)
    cam = Camera((0, 0, -8), (0, 0, 0, 1), 1.0, 1.0)
    n = 4
    z = torch.zeros
    with pytest.raises(ValueError):  # CPU tensors: the product path refuses, it never falls back
        render_splats(cam, (32, 32), z(n, 3), z(n, 2), z(n, 3), z(n, 4), z(n, 1, 3), z(n), False)


def test_scene_generator_deterministic():
    from brush_b200 import scene

    a = scene.make_scene(1000, 256, 256, 0, 1001)
    b = scene.make_scene(1000, 256, 256, 0, 1001)
    assert np.array_equal(a.means, b.means) and a.sh_coeffs.shape == (1000, 1, 3)
    assert np.allclose(np.linalg.norm(a.quats, axis=1), 1.0, atol=1e-5)


def test_rust_ffi_block_matches_the_header():
    """rust/brush-b200-sys/src/ffi.rs is generated from include/brush_b200.h (rust/gen_ffi.py): the committed file must
    be what the generator produces now, declare every prototype of the header with the same number of parameters, and
    the ctypes table of brush_b200/_lib.py must agree on the parameter counts too -- so neither binding can drift from
    the header the way a hand-written snippet once did."""
    import importlib.util

    spec = importlib.util.spec_from_file_location("gen_ffi", os.path.join(ROOT, "rust", "gen_ffi.py"))
    gen = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(gen)
    committed = open(os.path.join(ROOT, "rust", "brush-b200-sys", "src", "ffi.rs")).read()
    assert committed == gen.generate(), "ffi.rs is stale: run `python rust/gen_ffi.py`"
    hdr = open(os.path.join(ROOT, "include", "brush_b200.h")).read()
    _, _, structs, funcs = gen.parse(hdr)
    declared = set(re.findall(r"\b(brush_[a-z0-9_]+)\s*\(", gen.strip_comments(hdr)))
    assert {f[0] for f in funcs} == declared
    from brush_b200 import _lib

    for name, params, _ in funcs:
        m = re.search(r"pub fn " + name + r"\((.*?)\)", committed, flags=re.S)
        assert m is not None, name
        n_rust = len([p for p in m.group(1).split(",") if p.strip()])
        assert n_rust == len(params), name
        if name in _lib._SIGS:
            assert len(_lib._SIGS[name][1]) == len(params), f"ctypes signature of {name} disagrees with the header"
    assert set(_lib._SIGS) == declared  # the Python binding covers the whole ABI
    # struct sizes: the ctypes mirrors and the header-derived field lists agree on the field count
    import ctypes

    by_name = dict(structs)
    for cname in ("BrushUniforms", "BrushAux", "BrushAdamGroup", "BrushTrainState", "BrushViewExchange", "BrushDpSegment",
                  "BrushDpStep", "BrushSplatParams", "BrushRefineConfig", "BrushRefineCounts"):
        cls = getattr(_lib, cname)
        assert [f[0] for f in cls._fields_] == [f[0] for f in by_name[cname]], cname
        assert issubclass(cls, ctypes.Structure)
    # every entry point of the header has a safe wrapper in the sys crate (lib.rs: render path and primitives, train.rs:
    # the train step, dist.rs: the view-parallel step), called with as many arguments as the header declares
    crate = os.path.join(ROOT, "rust", "brush-b200-sys", "src")
    wrapped = {}
    for fname in ("lib.rs", "train.rs", "dist.rs"):
        src = open(os.path.join(crate, fname)).read()
        for m in re.finditer(r"ffi::(brush_[a-z0-9_]+)\(", src):
            i, depth = m.end(), 1
            j = i
            while depth:
                depth += {"(": 1, ")": -1}.get(src[j], 0)
                j += 1
            parts, d, cur = [], 0, ""
            for ch in src[i:j - 1]:
                d += {"(": 1, "[": 1, "{": 1, ")": -1, "]": -1, "}": -1}.get(ch, 0)
                if ch == "," and d == 0:
                    parts.append(cur)
                    cur = ""
                else:
                    cur += ch
            parts.append(cur)
            wrapped[m.group(1)] = len([x for x in parts if x.strip()])
    n_params = {name: len(params) for name, params, _ in funcs}
    assert set(wrapped) == declared, sorted(declared - set(wrapped))
    for name, n_args in wrapped.items():
        assert n_args == n_params[name], f"safe wrapper of {name} passes {n_args} arguments, the header declares {n_params[name]}"
    # the Burn-facing glue names every call it makes through the safe wrappers of the sys crate
    glue = open(os.path.join(ROOT, "rust", "brush-render-b200", "render_b200.rs")).read()
    sys_lib = open(os.path.join(ROOT, "rust", "brush-b200-sys", "src", "lib.rs")).read()
    for fn in ("render_forward", "render_backward"):
        assert f"ctx.{fn}(" in glue and f"ffi::brush_{fn}(" in sys_lib
