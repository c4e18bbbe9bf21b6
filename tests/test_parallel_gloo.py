"""world_size-2 gloo tests (CPU) of the view-parallel host logic: flat gradient layout, the single all-reduce,
global gradient scaling and view sharding."""
import os
import socket

import torch
import torch.distributed as dist
import torch.multiprocessing as mp


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from brush_b200.parallel import FlatGradLayout, allreduce_flat, global_grad_scale, view_for_rank

    n, k = 37, 16
    lay = FlatGradLayout(n, k)
    # every segment starts on a 16-byte boundary (the peer exchange moves float4s); payload = (13+3k+2)*n + loss
    assert all(o % 4 == 0 for o in lay.offsets) and sum(lay.sizes) == (13 + 3 * k + 2) * n + 4
    flat = lay.alloc("cpu")
    views = lay.views(flat)
    # each rank writes its own per-view gradients through the VIEWS (as brush_render_backward does via pointers)
    for i, v in enumerate(views):
        v.copy_(torch.full_like(v, float((rank + 1) * (i + 1))))
    allreduce_flat(flat, dist.group.WORLD)
    ok = True
    for i, v in enumerate(views):  # views alias the reduced buffer: sum over ranks of (rank+1)*(i+1)
        ok &= bool(torch.all(v == float(sum(r + 1 for r in range(world)) * (i + 1))))
    ok &= abs(global_grad_scale(1, world) - 1.0 / world) < 1e-12
    # view sharding: over 8 steps 2 ranks cover 16 views exactly once
    mine = [view_for_rank(s, world, rank, 16) for s in range(8)]
    gathered = [None] * world
    dist.all_gather_object(gathered, mine)
    ok &= sorted(sum(gathered, [])) == list(range(16))
    # peer-exchange ownership: emulate "reduce my slice, then hand it to everyone" with gloo point-to-point
    # collectives and check it equals the all-reduce -- the host-side contract of brush_dp_adam_step
    from brush_b200.parallel import owned_range

    torch.manual_seed(100 + rank)
    seg = torch.randn(4 * 25 + 3)
    full = seg.clone()
    dist.all_reduce(full)
    gathered_seg = [torch.empty_like(seg) for _ in range(world)]
    dist.all_gather(gathered_seg, seg)
    rebuilt = torch.zeros_like(seg)
    for r in range(world):
        b, e = owned_range(seg.numel(), world, r)
        rebuilt[b:e] = sum(g[b:e] for g in gathered_seg)  # what rank r would compute and store into every replica
    covered = owned_range(seg.numel(), world, world - 1)[1]
    ok &= covered == seg.numel() // 4 * 4 and bool(torch.allclose(rebuilt[:covered], full[:covered]))
    out[rank] = ok
    dist.destroy_process_group()


def test_flat_allreduce_and_view_sharding_world2():
    world = 2
    port = _free_port()
    mgr = mp.Manager()
    out = mgr.dict()
    mp.spawn(_worker, args=(world, port, out), nprocs=world, join=True)
    assert all(out[r] for r in range(world)) and len(out) == world


def test_single_rank_is_noop():
    from brush_b200.parallel import FlatGradLayout, allreduce_flat

    lay = FlatGradLayout(5, 1)
    flat = lay.alloc("cpu") + 3.0
    allreduce_flat(flat, None)
    assert torch.all(flat == 3.0) and lay.bytes_per_allreduce() == 4 * lay.total


def test_owned_ranges_partition_and_match_c_helper():
    import ctypes as C

    from brush_b200 import _lib
    from brush_b200.parallel import FlatGradLayout, FlatParamLayout, owned_range

    L = _lib.lib()  # dlopen only; no GPU call
    for count in (0, 4, 12, 100, 3_000_000, 48_000_000, 1_000_003 * 4):
        for world in (1, 2, 3, 8):
            prev = 0
            for rank in range(world):
                b, e = owned_range(count, world, rank)
                cb, ce = C.c_uint64(), C.c_uint64()
                L.brush_dp_owned_range(count, world, rank, C.byref(cb), C.byref(ce))
                assert (b, e) == (cb.value, ce.value) and b == prev and b % 4 == 0 and e % 4 == 0
                prev = e
            assert prev == count // 4 * 4
    g, p = FlatGradLayout(1001, 16), FlatParamLayout(1001, 16)
    for name, gname in p.GRAD_OF.items():  # every exchanged segment has the same padded length on both sides
        assert p.padded_size(name) == g.padded_size(gname) and p.offset(name) % 4 == 0 and g.offset(gname) % 4 == 0


def _two_view_inputs():
    """Seeded inputs shared by the gloo test below and tests/test_gpu_train.py::test_views_step_two_views_against_oracle."""
    import numpy as np

    from brush_b200 import Camera, scene
    from brush_b200.scene import bench_camera_params

    w, h, n, k = 112, 80, 1500, 16
    sc = scene.make_scene(n, w, h, 3, 21)
    pos, rot, fx, fy = bench_camera_params(w, h)
    cams = [Camera(pos, rot, fx, fy), Camera(np.asarray(pos) + np.array([0.4, -0.2, 0.1]), rot, fx * 1.1, fy * 1.1)]
    cam_args = [(c.world_to_local(), c.focal((w, h)), c.center((w, h))) for c in cams]
    rng = np.random.default_rng(3)
    v_outs = [(rng.standard_normal((h, w, 4)) * 0.01).astype(np.float32) for _ in cams]
    raw_rot = (sc.quats * 0.6).astype(np.float32)
    lrs = {"means": 1e-3, "log_scales": 5e-3, "rotation": 2e-3, "sh_coeffs": 3e-3, "raw_opacity": 2e-2}
    return sc, cams, cam_args, v_outs, raw_rot, lrs, k


def _oracle_step_worker(rank, world, port, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    import numpy as np

    import helpers as Hh
    from brush_b200.parallel import FlatGradLayout, allreduce_flat, global_grad_scale
    from oracle import pyoracle as O

    O.build()
    O.set_threads(1)
    sc, _, cam_args, v_outs, raw_rot, lrs, k = _two_view_inputs()
    n = sc.means.shape[0]
    # this rank's view only: the oracle is the render fwd + bwd (render.rs:55-323,468-625)
    g, norm, vis, (qhat, qn) = Hh.oracle_view_gradients(O, sc, *cam_args[rank], raw_rot, v_outs[rank])
    lay = FlatGradLayout(n, k)
    flat = lay.alloc("cpu")
    gm, gxy, gs, gq, gsh, go, gnorm, gvis, gloss = lay.views(flat)
    for dst, src in ((gm, g["means"]), (gs, g["log_scales"]), (gq, g["rotation"]), (gsh, g["sh_coeffs"]),
                     (go, g["raw_opacity"]), (gnorm, norm), (gvis, vis)):
        dst.copy_(torch.from_numpy(np.ascontiguousarray(src)))
    gloss[0] = 0.25 * (rank + 1)
    allreduce_flat(flat, dist.group.WORLD)  # the ONE exchange of the step
    scale = global_grad_scale(1, world)
    total = {"means": gm.numpy(), "log_scales": gs.numpy(), "sh_coeffs": gsh.numpy(), "raw_opacity": go.numpy(),
             "rotation": Hh.quat_norm_vjp(qhat, qn, gq.numpy()).astype(np.float32)}
    params = {"means": sc.means, "log_scales": sc.log_scales, "rotation": raw_rot, "sh_coeffs": sc.sh_coeffs,
              "raw_opacity": sc.raw_opacity}
    new_p, m, _ = Hh.adam_first_step(params, total, lrs, scale, k, 0.05)
    ok = True
    if rank == 0:
        ref = Hh.oracle_multi_view_step(O, sc, cam_args, raw_rot, v_outs, lrs, k, 0.05)
        for name in params:  # two addends: the all-reduce IS the in-order f32 sum, so the steps agree to the bit
            ok &= bool(np.array_equal(m[name], ref["m"][name])) and bool(np.array_equal(new_p[name], ref["params"][name]))
        ok &= bool(np.array_equal(gnorm.numpy() * scale, ref["grad_norm"])) and bool(np.array_equal(gvis.numpy(), ref["visible"]))
        ok &= abs(float(gloss[0]) * scale - 0.375) < 1e-7 and float(gvis.max()) == 2.0
    every = [None] * world
    dist.all_gather_object(every, float(np.abs(new_p["means"]).sum()))
    ok &= every[0] == every[1]  # identical replicas
    out[rank] = ok
    dist.destroy_process_group()


def test_two_view_step_through_the_oracle_world2():
    """Two gloo ranks, one view each: per-view ORACLE gradients and per-view refine statistics go through the flat
    layout and the single all-reduce, then Adam -- and must equal the one-process oracle step over both views in
    view order (tests/helpers.py::oracle_multi_view_step), which is what the GPU exchange kernels are compared with
    in tests/test_gpu_train.py::test_views_step_two_views_against_oracle."""
    import sys

    sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
    world, port = 2, _free_port()
    mgr = mp.Manager()
    out = mgr.dict()
    mp.spawn(_oracle_step_worker, args=(world, port, out), nprocs=world, join=True)
    assert len(out) == world and all(out[r] for r in range(world))
