"""TEST INFRASTRUCTURE ONLY -- CPU oracle for SplatTrainer::refine_splats.

numpy restatement of crates/brush-train/src/train.rs:395-553 (refine_splats), :140-177 (quaternion_vec_multiply),
:560-596 (prune_points), :598-621 (concat_splats), :205-209 (reset_opacity).  Only tests/ may import this module.

Parity status: UNPINNED -- the reference has no test or fixture for refinement; the restatement follows the source
line by line, including its quirk that the edits to the split splats themselves land on clones that are dropped
(train.rs:401-402 vs :519), so only the appended copies are smaller / offset.  The N(0, 0.5) draws are an input."""
from __future__ import annotations

import math

import numpy as np


def quaternion_vec_multiply(q: np.ndarray, v: np.ndarray) -> np.ndarray:  # train.rs:140-177
    qw, qx, qy, qz = (q[:, i:i + 1] for i in range(4))
    vx, vy, vz = (v[:, i:i + 1] for i in range(3))
    t1 = qw * vx + qy * vz - qz * vy
    t2 = qw * vy - qx * vz + qz * vx
    t3 = qw * vz + qx * vy - qy * vx
    t4 = qx * vx + qy * vy + qz * vz
    rx = vx + (qw * t1 + qx * t4 - qy * t3 + qz * t2) * 2.0
    ry = vy + (qw * t2 - qx * t3 + qy * t4 + qz * t1) * 2.0
    rz = vz + (qw * t3 + qx * t2 - qy * t1 + qz * t4) * 2.0
    return np.concatenate([rx, ry, rz], 1)


def refine(pre, post, grad_2d_accum, xy_grad_counts, samples, iter_before_increment, cfg):
    """pre / post: dicts of means, sh, rotation, raw_opacity, log_scales (float32 arrays) before / after the optimizer
    step.  samples: [num_split, 3] standard deviations already applied (N(0, 0.5) draws).  cfg: object with the
    TrainConfig fields.  Returns (dict of new parameters, (num_split, num_cloned, alpha_pruned, scale_pruned))."""
    f = np.float32
    grads = grad_2d_accum / np.maximum(xy_grad_counts, 1.0)                                   # :406
    big = grads >= f(cfg.densify_grad_thresh)                                                 # :408
    small = np.exp(post["log_scales"]).max(axis=1) < f(cfg.densify_size_thresh)               # :409-413
    clone_inds = np.nonzero(small & big)[0]                                                   # :421-428
    split_inds = np.nonzero(~small & big)[0]                                                  # :457-459
    names = ("means", "rotation", "sh", "raw_opacity", "log_scales")
    app = {n: [] for n in names}
    if len(clone_inds):                                                                       # :431-455 (pre-step values)
        for n in names:
            app[n].append(pre[n][clone_inds])
    if len(split_inds):                                                                       # :460-515
        cur_rot = post["rotation"][split_inds]
        cur_scale = np.exp(post["log_scales"][split_inds])
        app["rotation"].append(cur_rot)
        app["sh"].append(post["sh"][split_inds])
        app["raw_opacity"].append(post["raw_opacity"][split_inds])
        app["log_scales"].append(np.log(cur_scale / f(1.6)))
        app["means"].append(pre["means"][split_inds] + quaternion_vec_multiply(cur_rot, samples * cur_scale))
    new = {n: np.concatenate([post[n]] + app[n], 0) for n in names}                           # :519-536
    start = new["means"].shape[0]
    sig = 1.0 / (1.0 + np.exp(-new["raw_opacity"].astype(np.float64)))
    keep = ~(sig < cfg.cull_alpha_thresh)                                                     # :541-542
    new = {n: a[keep] for n, a in new.items()}
    alpha_pruned = start - new["means"].shape[0]
    keep = ~(np.exp(new["log_scales"]).max(axis=1) > f(cfg.cull_scale_thresh))                # :547-553
    new = {n: a[keep] for n, a in new.items()}
    scale_pruned = start - new["means"].shape[0]                                              # cumulative (:555)
    if (iter_before_increment // cfg.refine_every) % cfg.reset_alpha_every_refine == 0:       # :557-560
        v = math.log(cfg.reset_alpha_value / (1.0 - cfg.reset_alpha_value))                   # inverse_sigmoid
        new["raw_opacity"] = np.full_like(new["raw_opacity"], v)
    return new, (len(split_inds), len(clone_inds), alpha_pruned, scale_pruned)
