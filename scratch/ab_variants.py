"""A/B of the rasterizer variants (BRUSH_RASTER_FWD / BRUSH_RASTER_BWD) on the bench workload, one process.
usage: python scratch/ab_variants.py [steps]"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

import bench
from brush_b200 import Splats, get_context, scene
from brush_b200.train import SceneBatch, SplatTrainer, TrainConfig

steps = int(sys.argv[1]) if len(sys.argv) > 1 else 40
dev = torch.device("cuda", 0)
sc = scene.make_config_scene(bench.WORKLOAD, None)
W, H, N, K = sc.width, sc.height, sc.means.shape[0], sc.sh_coeffs.shape[1]
t = lambda a: torch.as_tensor(a, device=dev)
splats = Splats(t(sc.means), t(sc.sh_coeffs), t(sc.quats), t(sc.raw_opacity), t(sc.log_scales))
cams = bench.view_cameras(W, H)
ctx = get_context(dev)
gen = torch.Generator(device=dev)
gen.manual_seed(4242)
gt = torch.empty((bench.N_VIEWS, H, W, 3), dtype=torch.float32, device=dev)
with torch.no_grad():
    for v, cam in enumerate(cams):
        img, _ = splats.render(cam, (W, H), False, isect_capacity=bench.ISECT_CAPACITY)
        gt[v] = img[..., :3].clamp_(0.0, 1.0)
    splats.sh_coeffs.add_(torch.randn(splats.sh_coeffs.shape, device=dev, generator=gen) * 0.1)
    splats.raw_opacity.add_(torch.randn(splats.raw_opacity.shape, device=dev, generator=gen) * 0.3)
trainer = SplatTrainer(N, TrainConfig(max_refine_step=0), dev, sh_coeffs_per_channel=K,
                       isect_capacity=bench.ISECT_CAPACITY)
idx = [0]


def step():
    v = idx[0] % bench.N_VIEWS
    idx[0] += 1
    return trainer.step(SceneBatch(gt[v:v + 1], [cams[v]], 1.0), splats)[1]


for _ in range(10):
    step()
ctx.sync_check()
results = []
combos = [tuple(int(x) for x in c.split(":")) for c in os.environ.get("AB_COMBOS", "50:3").split(",")]
for fwd, bwd in combos:
    os.environ["BRUSH_RASTER_FWD"], os.environ["BRUSH_RASTER_BWD"] = str(fwd), str(bwd)
    for _ in range(3):
        step()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        step()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / steps
    ctx.set_timing(True)
    acc = {}
    for _ in range(5):
        step()
        for k_, v_ in ctx.stage_times_ms(torch.cuda.current_stream().cuda_stream).items():
            acc[k_] = acc.get(k_, 0.0) + v_ / 5
    ctx.set_timing(False)
    r = {"fwd": fwd, "bwd": bwd, "ms_step": round(ms, 4), "it_s": round(1000 / ms, 1),
         "Rasterize": round(acc["Rasterize"], 4), "RasterizeBackwards": round(acc["RasterizeBackwards"], 4)}
    print(json.dumps(r), flush=True)
    results.append(r)

if os.environ.get("BRUSH_DBG"):
    import ctypes as C
    from brush_b200 import _lib
    L = C.CDLL(_lib.SO_PATH)
    out = (C.c_ulonglong * 4)()
    L.brush_debug_counters(out, 1)
    st = step()
    L.brush_debug_counters(out, 1)
    print("dbg: survivors", out[0], "rows", out[1], "active pairs", out[2], "rows/survivors", out[1] / max(out[0], 1),
          "active lanes per row", out[2] / max(out[1], 1))
