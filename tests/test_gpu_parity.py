"""GPU parity tests: the CUDA path (through the C ABI) against the CPU oracle and the reference's golden vectors.

Bars (BASELINE.json north_star): intersection lists / tile offsets / sort order bit-exact (canonical form,
SURVEY.md 8a note vi); image within 1e-5 absolute; gradients within 1e-3 relative."""
import math

import numpy as np
import pytest
import torch

from helpers import all_close, crab, load_case, reference_camera

pytestmark = pytest.mark.gpu

DEV = "cuda:0"


def T(a, dtype=torch.float32):
    return torch.as_tensor(np.ascontiguousarray(a), device=DEV).to(dtype)


def record_counts(**kv):
    """Appends the exception counts a parity test OBSERVED (pixels over 1e-5, final_index mismatches, gradient entries
    outside 1e-3 relative) to gpurun_out/parity_counts.json, keyed by the running test; the file of the last full run
    is committed as profiles/r2_parity_counts.json and quoted in DESIGN.md section 4.  The asserts bound the counts;
    this records what they actually were."""
    import json
    import os

    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    path = os.environ.get("BRUSH_PARITY_COUNTS", os.path.join(root, "gpurun_out", "parity_counts.json"))
    test = os.environ.get("PYTEST_CURRENT_TEST", "?").split(" ")[0].split("::")[-1]
    try:
        os.makedirs(os.path.dirname(path), exist_ok=True)
        data = {}
        if os.path.exists(path):
            with open(path) as f:
                data = json.load(f)
        entry = data.setdefault(test, {})
        for k, v in kv.items():
            if isinstance(v, dict):
                entry.setdefault(k, {}).update(v)
            else:
                entry[k] = v
        with open(path, "w") as f:
            json.dump(data, f, indent=1, sort_keys=True)
    except OSError:
        pass


def grads_close(a, b, name):
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    scale = np.abs(b).max() + 1e-30
    bad = np.abs(a - b) > 1e-3 * np.abs(b) + 2e-5 * scale
    record_counts(grad_entries_outside_1e_3={name: [int(bad.sum()), int(bad.size)]})
    assert not bad.any(), f"{name}: {bad.sum()} / {bad.size} outside 1e-3 rel; max abs diff {np.abs(a - b).max():.3e}, scale {scale:.3e}"


# ------------------------------------------------------------------------------------------------ primitives
def test_sorting_reference_vectors(oracle):
    """brush-sort/src/lib.rs:164-217"""
    from brush_b200 import radix_argsort
    from test_oracle_golden import sort_vectors

    for keys in sort_vectors():
        vals = keys * 2 + 5
        n = torch.tensor([len(keys)], dtype=torch.int32, device=DEV)
        ko, vo = radix_argsort(T(keys.astype(np.int64), torch.int64).to(torch.int32), T(vals.astype(np.int64), torch.int64).to(torch.int32), n, 32)
        idx = np.argsort(keys, kind="stable")
        assert np.array_equal(ko.cpu().numpy().view(np.uint32), keys[idx])
        assert np.array_equal(vo.cpu().numpy().view(np.uint32), vals[idx])


@pytest.mark.parametrize("n,bits", [(1, 32), (255, 32), (4096, 32), (4097, 8), (100_003, 13), (1_200_000, 32),
                                    (3_000_000, 16)])
def test_sorting_big_vs_oracle(oracle, n, bits):
    """brush-sort/src/lib.rs:219-265 shape (clustered keys); payload = index, so stability is checked."""
    from brush_b200 import radix_argsort

    rng = np.random.default_rng(n)
    if bits == 32:
        keys = (rng.integers(0, 250, n) + np.arange(n) // 3).astype(np.uint32)
        keys[::7] = rng.integers(0, 2 ** 32, len(keys[::7]), dtype=np.uint64).astype(np.uint32)
    else:
        keys = rng.integers(0, 2 ** bits, n).astype(np.uint32)
    vals = np.arange(n, dtype=np.uint32)
    n_sort = max(1, n - n // 5)  # only the first n_sort entries take part; arrays are longer
    ko, vo = radix_argsort(T(keys.view(np.int32), torch.int32), T(vals.view(np.int32), torch.int32),
                           torch.tensor([n_sort], dtype=torch.int32, device=DEV), bits)
    rk, rv = oracle.radix_argsort_u32(keys, vals, n=n_sort, bits=bits)
    assert np.array_equal(ko.cpu().numpy().view(np.uint32)[:n_sort], rk[:n_sort])
    assert np.array_equal(vo.cpu().numpy().view(np.uint32)[:n_sort], rv[:n_sort])


def test_sorting_u64_fused_key(oracle):
    from brush_b200 import radix_argsort

    rng = np.random.default_rng(5)
    n = 700_001
    keys = (rng.integers(0, 8160, n).astype(np.uint64) << np.uint64(32)) | rng.integers(0, 2 ** 32, n).astype(np.uint64)
    keys[1000:2000] = keys[999]  # ties
    vals = np.arange(n, dtype=np.uint32)
    ko, vo = radix_argsort(T(keys.view(np.int64), torch.int64), T(vals.view(np.int32), torch.int32),
                           torch.tensor([n], dtype=torch.int32, device=DEV), 45)
    rk, rv = oracle.radix_argsort_u64(keys, vals, bits=45)
    assert np.array_equal(ko.cpu().numpy().view(np.uint64), rk)
    assert np.array_equal(vo.cpu().numpy().view(np.uint32), rv)


def test_prefix_sum_reference_vectors(oracle):
    """brush-prefix-sum/src/lib.rs:110-175"""
    from brush_b200 import prefix_sum

    cases = [np.array([1, 1, 1, 1]), 90 + np.arange(1024)]
    it = np.arange(512 * 16 + 123)
    cases.append(np.stack([2 + it, 0 * it, 32 + 0 * it, 512 + 0 * it, 30965 + 0 * it], 1).reshape(-1))
    rng = np.random.default_rng(1)
    cases.append(rng.integers(0, 40, 1_000_003))
    cases.append(rng.integers(0, 40, 1023))
    for d in cases:
        d = d.astype(np.uint32)
        got = prefix_sum(T(d.view(np.int32), torch.int32)).cpu().numpy().view(np.uint32)
        assert np.array_equal(got, oracle.prefix_sum_u32(d))


# ------------------------------------------------------------------------------------------------ golden vectors
@pytest.mark.parametrize("name", ["tiny_case", "basic_case"])
def test_reference_golden(name):
    """test_reference, crates/brush-render/src/render.rs:695-833, same tolerances."""
    from brush_b200 import Camera, Splats, fov_to_focal, focal_to_fov

    d = load_case(name)
    h, w, _ = d["out_img"].shape
    focal = fov_to_focal(math.pi * 0.5, w)
    cam = Camera((0.0, 0.0, -8.0), (0, 0, 0, 1), focal_to_fov(focal, w), focal_to_fov(focal, h))
    splats = Splats.from_safetensors(d, DEV)
    out, aux = splats.render(cam, (w, h), False)
    V = aux.read_num_visible()
    perm = aux.global_from_compact_gid[:V].cpu().numpy()
    proj = aux.projected_splats.cpu().numpy()
    assert all_close(proj[:V, 0:2], d["xys"][perm], 1e-4, 1e-10)
    assert all_close(proj[:V, 2:5], d["conics"][perm], 1e-4, 5e-7)
    out_rgb = out[..., :3]
    assert all_close(out_rgb.detach().cpu().numpy(), d["out_img"], 1e-4, 1e-9)
    loss = ((out_rgb - T(crab())) ** 2).mean()
    loss.backward()
    g = lambda t: t.grad.cpu().numpy()
    assert all_close(g(splats.xys_dummy), d["v_xy"], 1e-4, 1e-9)
    assert all_close(g(splats.raw_opacity), d["v_opacities"], 1e-4, 1e-10)
    assert all_close(g(splats.sh_coeffs), d["v_coeffs"], 1e-4, 1e-9)
    assert all_close(g(splats.log_scales), d["v_scales"], 1e-4, 1e-9)
    assert all_close(g(splats.means), d["v_means"], 1e-4, 1e-9)
    assert all_close(g(splats.rotation), d["v_quats"], 1e-1, 1e-1)


def test_renders_at_all():
    """render.rs:652-693"""
    from brush_b200 import Camera, render_splats

    cam = Camera((0, 0, 0), (0, 0, 0, 1), 0.5, 0.5)
    n = 8
    f = dict(device=DEV, dtype=torch.float32)
    quats = torch.tensor([[0.0, 0.0, 0.0, 1.0]], **f).repeat(n, 1)
    out, aux = render_splats(cam, (32, 32), torch.zeros(n, 3, **f).requires_grad_(), torch.zeros(n, 2, **f),
                             torch.ones(n, 3, **f) * 2.0, quats, torch.ones(n, 1, 3, **f), torch.zeros(n, **f), False)
    assert abs(float(out[..., :3].mean())) < 1e-5
    assert float(out[..., 3].mean()) == 0.0
    assert aux.read_num_visible() == 0 and aux.read_num_intersections() == 0


# ------------------------------------------------------------------------------------------------ oracle parity
def _run_both(oracle, sc, w, h, capacity=None, fused=False, v_out=None, seed=0):
    from brush_b200 import get_context, make_uniforms, render_backward, render_forward

    vm, focal, center = reference_camera(w, h)
    ou = oracle.make_uniforms(vm, focal, (w, h), center)
    ref_img, ref_aux = oracle.render_forward(ou, sc.means, sc.log_scales, sc.quats, sc.sh_coeffs, sc.raw_opacity,
                                             capacity=capacity)
    u = make_uniforms(vm, focal, (w, h), center)
    ctx = get_context(DEV)
    ctx.set_sort_mode(fused)
    tm, tl, tq, ts, to = T(sc.means), T(sc.log_scales), T(sc.quats), T(sc.sh_coeffs), T(sc.raw_opacity)
    img, aux = render_forward(u, tm, tl, tq, ts, to, False, isect_capacity=capacity, keep_tile_ids=True)
    ctx.set_sort_mode(False)
    grads = ref_grads = None
    if v_out is not None:
        grads = render_backward(aux, tm, tl, tq, ts.shape, to, img, T(v_out))
        ref_grads = oracle.render_backward(ref_aux, sc.means, sc.log_scales, sc.quats, sc.raw_opacity,
                                           sc.sh_coeffs.shape[1], ref_img, v_out)
    return img, aux, ref_img, ref_aux, grads, ref_grads


def _check_forward(img, aux, ref_img, ref_aux, fused=False):
    V, I = aux.read_num_visible(), aux.read_num_intersections()
    assert V == ref_aux.num_visible and I == ref_aux.num_intersections
    u = aux.uniforms_buffer.cpu().numpy()
    assert u[25] == V and u[26] == ref_aux.uniforms.total_splats and u[24] == ref_aux.uniforms.sh_degree
    gfc = aux.global_from_compact_gid.cpu().numpy().view(np.uint32)
    assert np.array_equal(gfc, ref_aux.global_from_compact_gid)           # depth order incl. zero tail
    assert np.array_equal(aux.cum_tiles_hit.cpu().numpy().view(np.uint32), ref_aux.cum_tiles_hit)
    assert np.array_equal(aux.tile_bins.cpu().numpy().view(np.uint32), ref_aux.tile_bins)
    cg = aux.compact_gid_from_isect.cpu().numpy().view(np.uint32)[:I]
    assert np.array_equal(cg, ref_aux.compact_gid_from_isect[:I])         # the intersection list itself
    if not fused:
        tid = aux.tile_id_from_isect.cpu().numpy().view(np.uint32)[:I]
        assert np.array_equal(tid, ref_aux.tile_id_from_isect[:I])
    proj = aux.projected_splats.cpu().numpy()[:V]
    rp = ref_aux.projected_splats[:V]
    assert np.array_equal(proj[:, [0, 1, 2, 3, 4, 8]], rp[:, [0, 1, 2, 3, 4, 8]])  # xy, conic, opac: bit-exact
    assert np.abs(proj[:, 5:8] - rp[:, 5:8]).max() < 2e-6                           # SH colour: float order only
    # image: 1e-5 absolute, except isolated pixels where some splat's alpha lies within float rounding of the 1/255
    # cut-off and exp() rounds differently on the two machines (the pair is then kept on one side only: an error of
    # at most alpha * T <= 1/255 on that pixel).  Those pixels are counted and bounded, never ignored.
    err = np.abs(img.cpu().numpy() - ref_img).max(axis=2)
    fi = aux.final_index.cpu().numpy().view(np.uint32)
    record_counts(pixels=int(err.size), pixels_over_1e_5=int((err > 1e-5).sum()), max_abs_err=float(err.max()),
                  final_index_mismatches=int((fi != ref_aux.final_index).sum()), num_visible=int(V), num_intersections=int(I))
    # bounds = a few times the counts observed on B200 (profiles/r2_parity_counts.json: at most 1 pixel per megapixel
    # over 1e-5 and 2 final_index mismatches in any case that goes through here)
    assert (err > 1e-5).sum() <= max(2, 4e-6 * err.size), f"{(err > 1e-5).sum()} pixels over 1e-5"
    assert err.max() < 1.0 / 255.0 + 1e-4 and np.median(err) < 1e-6
    assert (fi != ref_aux.final_index).sum() <= max(2, 1e-5 * err.size)  # decided by exp() rounding at the 1/255 threshold


@pytest.mark.parametrize("n,w,h,deg,seed", [(16_384, 256, 256, 0, 1001), (20_000, 333, 200, 3, 7),
                                            (3_000, 64, 48, 4, 9), (5_000, 130, 70, 1, 11), (5_000, 96, 96, 2, 13)])
def test_forward_backward_vs_oracle(oracle, n, w, h, deg, seed):
    from brush_b200 import scene

    sc = scene.make_scene(n, w, h, deg, seed)
    rng = np.random.default_rng(seed)
    target = rng.uniform(0, 1, (h, w, 3)).astype(np.float32)
    # MSE against a seeded random target, as render.rs:786-789; alpha gets a gradient too so v_out.a is exercised
    vm, focal, center = reference_camera(w, h)
    ou = oracle.make_uniforms(vm, focal, (w, h), center)
    ref_img, _ = oracle.render_forward(ou, sc.means, sc.log_scales, sc.quats, sc.sh_coeffs, sc.raw_opacity)
    v_out = np.zeros((h, w, 4), np.float32)
    v_out[..., :3] = 2.0 * (ref_img[..., :3] - target) / (h * w * 3)
    v_out[..., 3] = rng.uniform(-1, 1, (h, w)).astype(np.float32) / (h * w)
    img, aux, ref_img, ref_aux, grads, ref_grads = _run_both(oracle, sc, w, h, v_out=v_out)
    _check_forward(img, aux, ref_img, ref_aux)
    names = ["v_means", "v_xy", "v_log_scales", "v_quats", "v_coeffs", "v_raw_opac"]
    for gname, gt in zip(names, grads):
        grads_close(gt.cpu().numpy(), ref_grads[gname], gname)
    # dense gradients: exactly zero for non-visible splats (render.rs:571-575)
    V = aux.read_num_visible()
    vis = np.zeros(n, bool)
    vis[aux.global_from_compact_gid.cpu().numpy()[:V]] = True
    for gt in grads:
        assert not gt.cpu().numpy()[~vis].any()


def test_mix_case_shaped_scene_vs_oracle(oracle):
    """The INPUT recipe of the reference's third golden case, mix_case (NerfStudioRefGen.ipynb: 76 873 splats, means
    2000 * (U - 0.5), scales U * 15 + 0.05, raw opacities U(0, 1), SH degree 3, the 123 x 82 crab image, loss =
    mean((rgb - crab)^2), render.rs:786-789).  The fixture itself (gsplat's outputs) is missing from the checkout and
    its CUDA-RNG inputs cannot be regenerated bit for bit, so this is NOT a golden-vector test: the checker is the
    oracle, on a scene of the same shape -- a regime none of the bench scenes has (huge extent, most splats far away or
    behind the camera, the near ones covering the whole image)."""
    from brush_b200.scene import Scene

    from helpers import crab

    n, k = 76_873, 16
    rng = np.random.default_rng(4)
    f = np.float32
    means = (2000.0 * (rng.uniform(0, 1, (n, 3)) - 0.5)).astype(f)
    log_scales = np.log(rng.uniform(0, 1, (n, 3)) * 15.0 + 0.05).astype(f)
    coeffs = ((rng.uniform(0, 1, (n, k, 3)) - 0.5) * 0.5).astype(f)
    u, v, w_ = rng.uniform(0, 1, n), rng.uniform(0, 1, n), rng.uniform(0, 1, n)
    quats = np.stack([np.sqrt(1 - u) * np.sin(2 * np.pi * v), np.sqrt(1 - u) * np.cos(2 * np.pi * v),
                      np.sqrt(u) * np.sin(2 * np.pi * w_), np.sqrt(u) * np.cos(2 * np.pi * w_)], 1).astype(f)
    raw_opac = rng.uniform(0, 1, n).astype(f)
    target = crab()
    h, w = target.shape[:2]
    sc = Scene(means, log_scales, quats, coeffs, raw_opac, w, h, 3)
    vm, focal, center = reference_camera(w, h)
    ou = oracle.make_uniforms(vm, focal, (w, h), center)
    ref_img, _ = oracle.render_forward(ou, sc.means, sc.log_scales, sc.quats, sc.sh_coeffs, sc.raw_opacity)
    v_out = np.zeros((h, w, 4), f)
    v_out[..., :3] = 2.0 * (ref_img[..., :3] - target[..., :3]) / (h * w * 3)
    img, aux, ref_img, ref_aux, grads, ref_grads = _run_both(oracle, sc, w, h, v_out=v_out)
    assert 1000 < ref_aux.num_visible < n // 2 and ref_aux.num_intersections > ref_aux.num_visible  # 12 355 / 21 135
    _check_forward(img, aux, ref_img, ref_aux)
    for gname, gt in zip(["v_means", "v_xy", "v_log_scales", "v_quats", "v_coeffs", "v_raw_opac"], grads):
        grads_close(gt.cpu().numpy(), ref_grads[gname], gname)


def test_fused64_sort_mode_identical(oracle):
    from brush_b200 import scene

    sc = scene.make_scene(30_000, 400, 300, 0, 21)
    img, aux, ref_img, ref_aux, _, _ = _run_both(oracle, sc, 400, 300, fused=True)
    _check_forward(img, aux, ref_img, ref_aux, fused=True)


def test_many_splats_per_tile_and_depth_ties(oracle):
    """> 256 splats in one tile (multi-batch staging) and exact depth ties (order must fall back to gid)."""
    from brush_b200 import scene

    sc = scene.make_scene(6_000, 48, 32, 0, 33, sigma_px=(2.0, 6.0))
    sc.means[:, 2] = np.round(sc.means[:, 2] * 4) / 4  # many equal depths
    rng = np.random.default_rng(1)
    v_out = rng.uniform(-1, 1, (32, 48, 4)).astype(np.float32) / (32 * 48)
    img, aux, ref_img, ref_aux, grads, ref_grads = _run_both(oracle, sc, 48, 32, v_out=v_out)
    assert (ref_aux.tile_bins[..., 1] - ref_aux.tile_bins[..., 0]).max() > 256
    _check_forward(img, aux, ref_img, ref_aux)
    for gname, gt in zip(["v_means", "v_xy", "v_log_scales", "v_quats", "v_coeffs", "v_raw_opac"], grads):
        grads_close(gt.cpu().numpy(), ref_grads[gname], gname)


@pytest.mark.parametrize("fused", [False, True])
def test_large_splats_boxes_beyond_the_hit_mask(oracle, fused):
    """Splats whose tile boxes span 1 .. ~300 tiles: hit masks of more than 32 bits, boxes beyond 64 tiles (the
    per-tile fallback in map_intersects) and warps whose candidate lists need many passes.  Tile counts, the
    intersection list and tile_bins must stay bit-exact."""
    from brush_b200 import scene

    w, h = 400, 272
    sc = scene.make_scene(1_500, w, h, 1, 91, sigma_px=(1.0, 60.0))
    img, aux, ref_img, ref_aux, _, _ = _run_both(oracle, sc, w, h, fused=fused)
    hits = np.diff(np.concatenate([[0], ref_aux.cum_tiles_hit[:ref_aux.num_visible].astype(np.int64)]))
    assert hits.max() > 64 and (hits > 32).sum() > 10 and (hits <= 4).sum() > 10  # all three paths are exercised
    _check_forward(img, aux, ref_img, ref_aux, fused=fused)


def test_deferred_colours_forward_is_bit_identical():
    """brush_ctx_set_sh_ready_event / set_split_colors: the forward that evaluates SH colours in their own kernel just
    before the rasterizer (so that it can run under an in-flight SH exchange) must reproduce the fused forward bit for
    bit -- image, projected splats and every aux tensor -- with and without a pending event."""
    from brush_b200 import get_context, make_uniforms, render_forward, scene

    w, h = 240, 160
    sc = scene.make_scene(12_000, w, h, 3, 5)
    vm, focal, center = reference_camera(w, h)
    u = make_uniforms(vm, focal, (w, h), center)
    args = [T(sc.means), T(sc.log_scales), T(sc.quats), T(sc.sh_coeffs), T(sc.raw_opacity)]
    ctx = get_context(DEV)
    img0, aux0 = render_forward(u, *args, False)
    ctx.set_split_colors(True)
    try:
        img1, aux1 = render_forward(u, *args, False)
    finally:
        ctx.set_split_colors(False)
    # with an event: SH written on a side stream just before; the forward must pick the new values up
    side = torch.cuda.Stream()
    ev = torch.cuda.Event()
    sh2 = args[3].clone()
    torch.cuda.synchronize()
    with torch.cuda.stream(side):
        torch.cuda._sleep(20_000_000)  # ~10 ms: the forward's geometry stages finish long before this
        sh2.mul_(0.5)
        ev.record(side)
    ctx.set_sh_ready_event(ev.cuda_event)
    img2, aux2 = render_forward(u, args[0], args[1], args[2], sh2, args[4], False)
    torch.cuda.synchronize()
    img3, _ = render_forward(u, args[0], args[1], args[2], sh2, args[4], False)  # plain forward on the final values
    V, I = aux0.read_num_visible(), aux0.read_num_intersections()
    assert V == aux1.read_num_visible() and I == aux1.read_num_intersections() and V > 1000
    assert torch.equal(img0, img1) and torch.equal(aux0.projected_splats[:V], aux1.projected_splats[:V])
    assert torch.equal(aux0.final_index, aux1.final_index) and torch.equal(aux0.tile_bins, aux1.tile_bins)
    assert torch.equal(aux0.cum_tiles_hit, aux1.cum_tiles_hit)
    assert torch.equal(aux0.compact_gid_from_isect[:I], aux1.compact_gid_from_isect[:I])
    assert torch.equal(aux0.global_from_compact_gid, aux1.global_from_compact_gid)
    assert torch.equal(img2, img3) and not torch.equal(img2, img0)


def test_packed_forward_is_bit_identical():
    """BRUSH_RASTER_PACKED=1 selects the packed-FP32 compositing loop of the forward rasterizer (two survivors per
    FFMA2 / FMUL2 / FADD2; an opt-in A/B, profiles/r2p_rasterizer_experiments.md).  It performs the scalar loop's
    operations in the scalar loop's order, so image and final_index must be the same bits -- f32 and packed-u32 output,
    many batches per tile included.  The switch is read once per process: two child processes hash their outputs."""
    import os
    import subprocess
    import sys

    code = r"""
import hashlib, sys
import numpy as np, torch
sys.path.insert(0, sys.argv[1]); sys.path.insert(0, sys.argv[1] + "/tests")
from helpers import reference_camera
from brush_b200 import make_uniforms, render_forward, scene
h = lambda x: hashlib.sha256(x.contiguous().cpu().numpy().tobytes()).hexdigest()
for n, w, hh, deg, seed in ((30_000, 320, 200, 3, 21), (6_000, 48, 32, 1, 22)):
    sc = scene.make_scene(n, w, hh, deg, seed)
    vm, focal, center = reference_camera(w, hh)
    u = make_uniforms(vm, focal, (w, hh), center)
    t = lambda a: torch.as_tensor(np.ascontiguousarray(a), device="cuda:0")
    args = [t(sc.means), t(sc.log_scales), t(sc.quats), t(sc.sh_coeffs), t(sc.raw_opacity)]
    img, aux = render_forward(u, *args, False)
    img8, _ = render_forward(u, *args, True)
    torch.cuda.synchronize()
    print(n, aux.read_num_intersections(), h(img), h(aux.final_index), h(img8))
"""
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    outs = []
    for packed in ("0", "1"):
        env = dict(os.environ, BRUSH_RASTER_PACKED=packed)
        r = subprocess.run([sys.executable, "-c", code, root], env=env, capture_output=True, text=True, timeout=300)
        assert r.returncode == 0, r.stderr[-2000:]
        outs.append(r.stdout.strip().splitlines())
    assert len(outs[0]) == 2 and outs[0] == outs[1], (outs[0], outs[1])
    assert int(outs[0][1].split()[1]) > 48 * 32 // 256 * 600  # the small image holds several batches per tile


def test_raster_u32(oracle):
    """Packed RGBA8 output (rasterize.wgsl:106-109): equal up to 1 LSB per channel (exp rounding)."""
    from brush_b200 import make_uniforms, render_forward, scene

    w, h = 200, 120
    sc = scene.make_scene(8_000, w, h, 0, 5)
    vm, focal, center = reference_camera(w, h)
    ref, _ = oracle.render_forward(oracle.make_uniforms(vm, focal, (w, h), center), sc.means, sc.log_scales, sc.quats,
                                   sc.sh_coeffs, sc.raw_opacity, raster_u32=True)
    out, _ = render_forward(make_uniforms(vm, focal, (w, h), center), T(sc.means), T(sc.log_scales), T(sc.quats),
                            T(sc.sh_coeffs), T(sc.raw_opacity), True)
    got = out.cpu().numpy().view(np.uint32).reshape(h, w)
    for sh in (0, 8, 16, 24):
        a, b = (got >> sh) & 255, (ref >> sh) & 255
        assert np.abs(a.astype(int) - b.astype(int)).max() <= 1
    assert (got != ref).mean() < 0.02


def test_capacity_overflow_is_reported(oracle):
    from brush_b200 import get_context, make_uniforms, render_forward, scene
    from brush_b200._lib import CapacityError

    w, h = 128, 128
    sc = scene.make_scene(4_000, w, h, 0, 3)
    vm, focal, center = reference_camera(w, h)
    u = make_uniforms(vm, focal, (w, h), center)
    args = [T(sc.means), T(sc.log_scales), T(sc.quats), T(sc.sh_coeffs), T(sc.raw_opacity)]
    _, aux = render_forward(u, *args, False, isect_capacity=1000)
    with pytest.raises(CapacityError):
        get_context(DEV).sync_check()
    assert aux.read_num_intersections() == 1000
    # same truncation as the oracle / map_gaussian_to_intersects.wgsl:40
    ref_img, ref_aux = oracle.render_forward(oracle.make_uniforms(vm, focal, (w, h), center), sc.means, sc.log_scales,
                                             sc.quats, sc.sh_coeffs, sc.raw_opacity, capacity=1000)
    assert ref_aux.status == -2
    assert np.array_equal(aux.compact_gid_from_isect.cpu().numpy().view(np.uint32)[:1000],
                          ref_aux.compact_gid_from_isect[:1000])
    _, aux = render_forward(u, *args, False)  # default capacity is fine again
    get_context(DEV).sync_check()


def test_empty_and_single():
    from brush_b200 import make_uniforms, render_forward

    vm, focal, center = reference_camera(40, 24)
    u = make_uniforms(vm, focal, (40, 24), center)
    f = dict(device=DEV, dtype=torch.float32)
    out, aux = render_forward(u, torch.zeros(0, 3, **f), torch.zeros(0, 3, **f), torch.zeros(0, 4, **f),
                              torch.zeros(0, 1, 3, **f), torch.zeros(0, **f), False)
    assert float(out.abs().max()) == 0.0 and aux.read_num_visible() == 0
    q = torch.tensor([[1.0, 0, 0, 0]], **f)
    out, aux = render_forward(u, torch.zeros(1, 3, **f), torch.zeros(1, 3, **f), q, torch.ones(1, 1, 3, **f),
                              torch.ones(1, **f) * 2, False)
    assert aux.read_num_visible() == 1 and float(out[..., 3].max()) > 0.5


def test_full_size_config4_vs_oracle(oracle):
    """BASELINE.json configs[3] at full size: 1M Gaussians, SH3, 1920x1080.  Integer outputs bit-exact; image within
    1e-5 except for the (counted) pixels where a splat sits exactly on the alpha = 1/255 threshold and exp() rounding
    decides differently on the two machines; gradients within 1e-3 relative."""
    from brush_b200 import scene

    oracle.set_threads(oracle.max_threads())
    try:
        sc = scene.make_config_scene("cfg4_1m_1080p")
        w, h = sc.width, sc.height
        v_out = np.full((h, w, 4), 1.0 / (4 * h * w), np.float32)
        img, aux, ref_img, ref_aux, grads, ref_grads = _run_both(oracle, sc, w, h, capacity=16 << 20, v_out=v_out)
    finally:
        oracle.set_threads(1)
    V, I = aux.read_num_visible(), aux.read_num_intersections()
    assert V == ref_aux.num_visible and I == ref_aux.num_intersections
    assert np.array_equal(aux.global_from_compact_gid.cpu().numpy().view(np.uint32), ref_aux.global_from_compact_gid)
    assert np.array_equal(aux.cum_tiles_hit.cpu().numpy().view(np.uint32), ref_aux.cum_tiles_hit)
    assert np.array_equal(aux.tile_bins.cpu().numpy().view(np.uint32), ref_aux.tile_bins)
    assert np.array_equal(aux.compact_gid_from_isect.cpu().numpy().view(np.uint32)[:I], ref_aux.compact_gid_from_isect[:I])
    assert np.array_equal(aux.tile_id_from_isect.cpu().numpy().view(np.uint32)[:I], ref_aux.tile_id_from_isect[:I])
    err = np.abs(img.cpu().numpy() - ref_img).max(axis=2)
    flips = int((err > 1e-5).sum())
    print(f"full size: V={V} I={I} max err {err.max():.2e}, pixels over 1e-5: {flips} of {err.size}")
    fi = aux.final_index.cpu().numpy().view(np.uint32)
    record_counts(pixels=int(err.size), pixels_over_1e_5=flips, max_abs_err=float(err.max()),
                  final_index_mismatches=int((fi != ref_aux.final_index).sum()), num_visible=int(V), num_intersections=int(I))
    # observed on B200: 7 pixels of 2 073 600 over 1e-5, 4 final_index mismatches (profiles/r2_parity_counts.json)
    assert flips <= 16 and err.max() < 1.0 / 255.0 + 1e-4 and int((fi != ref_aux.final_index).sum()) <= 12
    for gname, gt in zip(["v_means", "v_xy", "v_log_scales", "v_quats", "v_coeffs", "v_raw_opac"], grads):
        a, b = gt.cpu().numpy().astype(np.float64), ref_grads[gname].astype(np.float64)
        bad = np.abs(a - b) > 1e-3 * np.abs(b) + 2e-5 * np.abs(b).max()
        record_counts(grad_entries_outside_1e_3={gname: [int(bad.sum()), int(bad.size)]})
        assert bad.sum() <= 32, f"{gname}: {bad.sum()} of {bad.size}"  # observed: at most 10 entries (v_xy)


@pytest.mark.parametrize("name", ["cfg2_100k_1024", "cfg3_300k_800"])
def test_baseline_configs_2_and_3_vs_oracle(oracle, name):
    """BASELINE.json configs[1] and [2] at full size (100k Gaussians at 1024^2, SH 0; 300k at 800^2, SH 3)."""
    from brush_b200 import scene

    oracle.set_threads(oracle.max_threads())
    try:
        sc = scene.make_config_scene(name)
        w, h = sc.width, sc.height
        rng = np.random.default_rng(17)
        v_out = (rng.uniform(-1, 1, (h, w, 4)) / (h * w)).astype(np.float32)
        img, aux, ref_img, ref_aux, grads, ref_grads = _run_both(oracle, sc, w, h, v_out=v_out)
    finally:
        oracle.set_threads(1)
    _check_forward(img, aux, ref_img, ref_aux)
    for gname, gt in zip(["v_means", "v_xy", "v_log_scales", "v_quats", "v_coeffs", "v_raw_opac"], grads):
        a, b = gt.cpu().numpy().astype(np.float64), ref_grads[gname].astype(np.float64)
        bad = np.abs(a - b) > 1e-3 * np.abs(b) + 2e-5 * np.abs(b).max()
        # a pair sitting on the alpha = 1/255 threshold is kept by one machine only (see _check_forward): the splats
        # of those few pairs see a different gradient; counted and bounded
        record_counts(grad_entries_outside_1e_3={gname: [int(bad.sum()), int(bad.size)]})
        assert bad.sum() <= 32, f"{gname}: {bad.sum()} of {bad.size}"  # observed: at most 9 entries (v_quats, cfg 2)


@pytest.mark.parametrize("fused", [False, True])
def test_config5_6m_4k_digest_vs_oracle(fused):
    """BASELINE.json configs[4] PINNED: the oracle was run once offline at full size (tests/golden/make_cfg5_digest.py,
    minutes of CPU) and its integer outputs committed as SHA-256 digests (tests/golden/cfg5_digest.json).  The GPU
    lists -- depth order, scan, tile bins and the ~77 M-entry tile-sorted intersection list -- must hash to the same
    values: bit-exact keys AND values at 1e8 sort elements, as crates/brush-sort/src/lib.rs:219-265 demands, in both
    tile-sort modes."""
    import json
    import os
    import sys

    from brush_b200 import get_context, make_uniforms, render_forward, scene

    gold = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
    sys.path.insert(0, gold)
    from make_cfg5_digest import CAPACITY, WORKLOAD, digest

    with open(os.path.join(gold, "cfg5_digest.json")) as f:
        want = json.load(f)
    sc = scene.make_config_scene(WORKLOAD)
    w, h = sc.width, sc.height
    vm, focal, center = reference_camera(w, h)
    u = make_uniforms(vm, focal, (w, h), center)
    args = [T(sc.means), T(sc.log_scales), T(sc.quats), T(sc.sh_coeffs), T(sc.raw_opacity)]
    ctx = get_context(DEV)
    ctx.set_sort_mode(fused)
    try:
        _, aux = render_forward(u, *args, False, isect_capacity=CAPACITY)
        ctx.sync_check()
    finally:
        ctx.set_sort_mode(False)
    got = digest(aux.read_num_visible(), aux.read_num_intersections(), aux.global_from_compact_gid.cpu().numpy(),
                 aux.cum_tiles_hit.cpu().numpy(), aux.tile_bins.cpu().numpy(), aux.compact_gid_from_isect.cpu().numpy())
    for k, v in got.items():
        assert want[k] == v, k


def test_config5_6m_4k_properties():
    """BASELINE.json configs[4]: 6M Gaussians, SH 3, 3840x2160, ~10^8 (tile, depth) keys.  Besides the digest
    comparison above, the size-independent properties of the path are checked: the scan total is
    the intersection count, the intersection list is sorted by tile and, inside a tile, by compact (= depth) order,
    tile_bins are exactly the run boundaries of the sorted tile ids, the compact order is (depth, gid) ascending,
    and a second forward reproduces every output bit for bit."""
    from brush_b200 import get_context, make_uniforms, render_forward, scene

    sc = scene.make_config_scene("cfg5_6m_4k")
    w, h, n = sc.width, sc.height, sc.means.shape[0]
    vm, focal, center = reference_camera(w, h)
    u = make_uniforms(vm, focal, (w, h), center)
    args = [T(sc.means), T(sc.log_scales), T(sc.quats), T(sc.sh_coeffs), T(sc.raw_opacity)]
    cap = 160 << 20
    img, aux = render_forward(u, *args, False, isect_capacity=cap, keep_tile_ids=True)
    get_context(DEV).sync_check()  # raises on intersection-capacity overflow
    V, I = aux.read_num_visible(), aux.read_num_intersections()
    assert 0 < V <= n and 50_000_000 < I < cap, (V, I)
    cum = aux.cum_tiles_hit[:V].long()
    assert int(cum[-1]) == I and bool((cum[1:] >= cum[:-1]).all())
    tid = aux.tile_id_from_isect[:I].long()
    cg = aux.compact_gid_from_isect[:I].long()
    assert bool((tid[1:] >= tid[:-1]).all())                                   # sorted by tile
    same = tid[1:] == tid[:-1]
    assert bool((cg[1:][same] > cg[:-1][same]).all())                          # stable: depth order inside a tile
    assert int(cg.max()) < V
    tiles = ((w + 15) // 16) * ((h + 15) // 16)
    counts = torch.bincount(tid, minlength=tiles)
    ends = torch.cumsum(counts, 0)
    bins = aux.tile_bins.view(-1, 2).long()
    nz = counts > 0
    assert bool((bins[nz, 1] == ends[nz]).all()) and bool((bins[nz, 0] == (ends - counts)[nz]).all())
    assert bool(((bins[~nz, 1] - bins[~nz, 0]) == 0).all())
    # compact order = ascending (depth, gid)
    gfc = aux.global_from_compact_gid[:V].long()
    vmt = torch.as_tensor(np.asarray(vm, np.float32), device=DEV)  # row-major world-to-camera
    depth = args[0][gfc] @ vmt[2, :3] + vmt[2, 3]
    d = depth[1:] - depth[:-1]
    assert float(d.min()) > -1e-4                                               # float re-evaluation noise only
    assert int(torch.unique(gfc).numel()) == V
    out = img.float()
    # (colours are SH + 0.5 unclamped, so only alpha is bounded)
    assert bool(torch.isfinite(out).all()) and float(out[..., 3].max()) <= 1.0 + 1e-6 and float(out[..., 3].min()) >= 0.0
    # determinism: every output of a second forward is identical
    img2, aux2 = render_forward(u, *args, False, isect_capacity=cap, keep_tile_ids=True)
    assert torch.equal(img, img2) and torch.equal(aux.final_index, aux2.final_index)
    assert torch.equal(aux.compact_gid_from_isect[:I], aux2.compact_gid_from_isect[:I])
    assert torch.equal(aux.tile_bins, aux2.tile_bins) and torch.equal(aux.global_from_compact_gid, aux2.global_from_compact_gid)
