"""Host time to enqueue one training step (is the e2e loop CPU-bound?)."""
import sys, time, torch, numpy as np
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
gt_host = gt.cpu().pin_memory()
tr = SplatTrainer(N, TrainConfig(max_refine_step=0), dev, sh_coeffs_per_channel=K, isect_capacity=16_777_216)
for s in range(10):
    splats, _ = tr.step(SceneBatch(gt[s % 16:s % 16 + 1], [cams[s % 16]], 1.0), splats)
torch.cuda.synchronize()
t0 = time.perf_counter()
for s in range(50):
    splats, _ = tr.step(SceneBatch(gt[s % 16:s % 16 + 1], [cams[s % 16]], 1.0), splats)
t1 = time.perf_counter()
torch.cuda.synchronize()
t2 = time.perf_counter()
print(f"host enqueue {1e3 * (t1 - t0) / 50:.3f} ms/step, total {1e3 * (t2 - t0) / 50:.3f} ms/step")
torch.cuda.synchronize()
t0 = time.perf_counter()
for s in range(50):
    b = tr.prefetch(SceneBatch(gt_host[s % 16:s % 16 + 1], [cams[s % 16]], 1.0))
    splats, _ = tr.step(b, splats)
t1 = time.perf_counter()
torch.cuda.synchronize()
t2 = time.perf_counter()
print(f"with prefetch: host enqueue {1e3 * (t1 - t0) / 50:.3f} ms/step, total {1e3 * (t2 - t0) / 50:.3f} ms/step")
import cProfile, pstats
pr = cProfile.Profile()
pr.enable()
for s in range(50):
    splats, _ = tr.step(SceneBatch(gt[s % 16:s % 16 + 1], [cams[s % 16]], 1.0), splats)
pr.disable()
torch.cuda.synchronize()
pstats.Stats(pr).sort_stats("cumulative").print_stats(18)
