"""views_adam_kernel with world = 1..8 emulated on ONE GPU (all row blocks local): the kernel's cost without NVLink."""
import ctypes as C, sys, torch, numpy as np
sys.path.insert(0, ".")
import bench
from brush_b200 import _lib, get_context, scene
from brush_b200.render import make_uniforms, render_forward

dev = torch.device("cuda", 0)
L = _lib.lib(); ctx = get_context(dev)
N, K, W, H = 1_000_000, 16, 1920, 1080
sc = scene.make_scene(N, W, H, 3, 4)
t = lambda a: torch.as_tensor(a, device=dev)
params = {"means": t(sc.means), "log_scales": t(sc.log_scales), "rotation": t(sc.quats), "sh_coeffs": t(sc.sh_coeffs), "raw_opacity": t(sc.raw_opacity)}
norm_rot = params["rotation"].clone()
cams = bench.view_cameras(W, H)
block = N * 12 + 32
for world in (1, 2, 4, 8):
    rows = torch.zeros(world * block, device=dev)
    flags = torch.zeros(64, dtype=torch.int32, device=dev)
    v_out = torch.randn((H, W, 4), device=dev) * 1e-6
    def xch(r, epoch):
        x = _lib.BrushViewExchange(); x.world, x.rank, x.epoch = world, r, epoch
        for i in range(world):
            x.rows[i], x.flags[i] = rows.data_ptr() + 4 * block * i, flags.data_ptr()
        return x
    for r in range(world):
        cam = cams[r]
        u = make_uniforms(cam.world_to_local(), cam.focal((W, H)), (W, H), cam.center((W, H)))
        pred, aux = render_forward(u, params["means"], params["log_scales"], norm_rot, params["sh_coeffs"], params["raw_opacity"], False, isect_capacity=16_777_216, ctx=ctx)
        caux = aux._c(); x = xch(r, 1)
        ctx.check(L.brush_render_backward_rows(ctx.handle, N, K, pred.data_ptr(), v_out.data_ptr(), C.byref(caux), None, C.byref(x), None))
    S = _lib.BrushTrainState(); state = {}
    for name, p in params.items():
        state[name] = (torch.zeros_like(p), torch.zeros_like(p))
        setattr(S, name, p.data_ptr()); setattr(S, "m_" + name, state[name][0].data_ptr()); setattr(S, "v_" + name, state[name][1].data_ptr())
    for kk, v in dict(lr_mean=1e-7, lr_scale=1e-7, lr_rotation=1e-7, lr_coeffs_dc=1e-7, lr_opac=1e-7, sh_rest_lr_scale=0.05).items():
        setattr(S, kk, v)
    x = xch(0, 1)
    acc = torch.zeros(2, N, device=dev)
    def run(step):
        ctx.check(L.brush_views_adam_step(ctx.handle, N, K, C.byref(S), norm_rot.data_ptr(), C.byref(x), acc[0].data_ptr(), acc[1].data_ptr(), None, 0.9, 0.999, 1e-15, step, 1.0 / world, None))
    for i in range(3): run(i + 1)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(20): run(i + 4)
    e1.record(); torch.cuda.synchronize()
    print(f"world {world}: views_adam {e0.elapsed_time(e1) / 20 * 1e3:.1f} us (all rows local)")
