"""Per-view cost of the bench views before / after N optimizer steps (is multi-GPU skew view imbalance or drift?)."""
import sys, torch, numpy as np
sys.path.insert(0, ".")
import bench
from brush_b200 import Splats, get_context, scene
from brush_b200.train import SceneBatch, SplatTrainer, TrainConfig

dev = torch.device("cuda", 0)
N, K, W, H = 1_000_000, 16, 1920, 1080
sc = scene.make_scene(N, W, H, 3, 4)
t = lambda a: torch.as_tensor(a, device=dev)
splats = Splats(t(sc.means), t(sc.sh_coeffs), t(sc.quats), t(sc.raw_opacity), t(sc.log_scales))
cams = bench.view_cameras(W, H)
gt = torch.rand((16, H, W, 3), device=dev)
ctx = get_context(dev)
tr = SplatTrainer(N, TrainConfig(max_refine_step=0), dev, sh_coeffs_per_channel=K, isect_capacity=24_000_000)

def per_view(tag):
    trn = SplatTrainer(N, TrainConfig(max_refine_step=0), dev, sh_coeffs_per_channel=K, isect_capacity=24_000_000, optimize=False)
    ctx.set_timing(True)
    out = []
    for v in range(16):
        for _ in range(2):
            trn.step(SceneBatch(gt[v:v+1], [cams[v]], 1.0), splats)
        tm = ctx.stage_times_ms(torch.cuda.current_stream().cuda_stream)
        out.append(round(tm["Rasterize"] + tm["RasterizeBackwards"], 3))
    ctx.set_timing(False)
    print(tag, out, "mean", round(float(np.mean(out)), 3), "max", max(out))

per_view("start ")
for s in range(70):
    splats, _ = tr.step(SceneBatch(gt[s % 16:s % 16 + 1], [cams[s % 16]], 1.0), splats)
per_view("step70")
for s in range(70, 300):
    splats, _ = tr.step(SceneBatch(gt[s % 16:s % 16 + 1], [cams[s % 16]], 1.0), splats)
per_view("step300")
