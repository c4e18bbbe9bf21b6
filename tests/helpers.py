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


# ---- an ORACLE-backed view-parallel training step ---------------------------------------------------
# The reference trains one view per step on one device (train.rs:216-219); its batch semantics are "loss = mean over
# the batch" (train.rs:229-268) and "refine statistics per view, before any reduction" (train.rs:300-314).  These
# helpers restate one optimizer step over several views with the CPU oracle and numpy, so that the multi-GPU exchange
# kernels (views / peer / nccl) are checked against the oracle and not only against each other.

def oracle_view_gradients(O, sc, viewmat, focal, center, raw_rot, v_out):
    """One view's dense gradients w.r.t. (means, log_scales, RAW rotation, sh, raw opacity) through the oracle:
    normalise rotation -> render_forward -> render_backward(v_out) -> normalisation VJP; plus the view's refine
    statistics (|v_xy * (w/2, h/2)| and the visibility flag per splat, train.rs:300-314)."""
    w, h = sc.width, sc.height
    qhat, qn = normalize_quats(raw_rot)
    u = O.make_uniforms(viewmat, focal, (w, h), center)
    img, aux = O.render_forward(u, sc.means, sc.log_scales, qhat, sc.sh_coeffs, sc.raw_opacity)
    g = O.render_backward(aux, sc.means, sc.log_scales, qhat, sc.raw_opacity, sc.sh_coeffs.shape[1], img, v_out)
    n = sc.means.shape[0]
    vis = np.zeros(n, np.float32)
    vis[aux.global_from_compact_gid[:aux.num_visible]] = 1.0
    xy = g["v_xy"] * np.array([0.5 * w, 0.5 * h], np.float32)
    norm = np.sqrt((xy * xy).sum(1)).astype(np.float32)
    grads = {"means": g["v_means"], "log_scales": g["v_log_scales"], "rotation": g["v_quats"], "sh_coeffs": g["v_coeffs"],
             "raw_opacity": g["v_raw_opac"]}
    return grads, norm, vis, (qhat, qn)


def adam_first_step(params, grads, lrs, grad_scale, k, sh_rest_lr_scale, b1=0.9, b2=0.999, eps=1e-15):
    """Burn Adam at t = 1 from zero moments (train.rs:318-359) in float64; returns (new params, m, v)."""
    out, ms, vs = {}, {}, {}
    for name, p in params.items():
        g = grads[name].astype(np.float64) * grad_scale
        m, v = (1 - b1) * g, (1 - b2) * g * g
        upd = (m / (1 - b1)) / (np.sqrt(v / (1 - b2)) + eps)
        lr = np.full_like(upd, lrs[name])
        if name == "sh_coeffs" and k > 1:
            lr[:, 1:, :] *= sh_rest_lr_scale  # the 1/lr_coeffs_sh_scale lerp of train.rs:334-348
        out[name], ms[name], vs[name] = p.astype(np.float64) - lr * upd, m, v
    return out, ms, vs


def oracle_multi_view_step(O, sc, cams, raw_rot, v_outs, lrs, k, sh_rest_lr_scale):
    """The whole step on one process: per-view oracle gradients summed IN VIEW ORDER (f32, as the exchange kernels
    do), statistics summed per view, rotation VJP applied to the summed gradient (it is linear), Adam with
    grad_scale = 1/views.  cams: list of (viewmat, focal, center)."""
    total, norm_sum, vis_sum, qs = None, 0.0, 0.0, None
    for (vm, focal, center), v_out in zip(cams, v_outs):
        g, norm, vis, qs = oracle_view_gradients(O, sc, vm, focal, center, raw_rot, v_out)
        total = g if total is None else {kk: total[kk] + g[kk] for kk in g}
        norm_sum = norm_sum + norm
        vis_sum = vis_sum + vis
    scale = 1.0 / len(cams)
    total["rotation"] = quat_norm_vjp(qs[0], qs[1], total["rotation"]).astype(np.float32)
    params = {"means": sc.means, "log_scales": sc.log_scales, "rotation": raw_rot, "sh_coeffs": sc.sh_coeffs,
              "raw_opacity": sc.raw_opacity}
    new_p, m, v = adam_first_step(params, total, lrs, scale, k, sh_rest_lr_scale)
    return {"grads": total, "params": new_p, "m": m, "v": v, "grad_norm": norm_sum * scale, "visible": vis_sum}
