"""torchrun: event timeline of the overlapped peer exchange (where does the SH exchange sit relative to the next forward?)."""
import os, sys
sys.path.insert(0, '.')
import numpy as np
import torch, torch.distributed as dist
rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(rank); dev = torch.device("cuda", rank)
dist.init_process_group("nccl", device_id=dev)
import bench
from brush_b200 import Splats, scene
import brush_b200.train as T
from brush_b200.parallel import view_for_rank
sc = scene.make_config_scene("cfg4_1m_1080p", None)
W, H, N, K = sc.width, sc.height, sc.means.shape[0], sc.sh_coeffs.shape[1]
t = lambda a: torch.as_tensor(a, device=dev)
splats = Splats(t(sc.means), t(sc.sh_coeffs), t(sc.quats), t(sc.raw_opacity), t(sc.log_scales))
cams = bench.view_cameras(W, H)
gt = torch.rand((bench.N_VIEWS, H, W, 3), device=dev)
tr = T.SplatTrainer(N, T.TrainConfig(max_refine_step=0, warmup_steps=0), dev, sh_coeffs_per_channel=K,
                    process_group=dist.group.WORLD, isect_capacity=16777216, exchange="peer")
splats = tr.adopt(splats)
E = lambda: torch.cuda.Event(enable_timing=True)
log = []
cur = {}
orig_fwd, orig_peer = T.render_forward, tr._peer_step
def fwd(*a, **k):
    cur["f0"] = E(); cur["f0"].record()
    r = orig_fwd(*a, **k)
    cur["f1"] = E(); cur["f1"].record()
    return r
def peer(*a, **k):
    cur["x0"] = E(); cur["x0"].record()
    r = orig_peer(*a, **k)
    cur["x1"] = E(); cur["x1"].record()
    if tr._xchg_stream is not None and tr.overlap_exchange:
        cur["sh1"] = E(); cur["sh1"].record(tr._xchg_stream)
    return r
T.render_forward = fwd
tr._peer_step = peer
steps = 14
for s in range(steps):
    cur = {"s": E()}
    cur["s"].record()
    v = view_for_rank(s, world, rank, bench.N_VIEWS)
    splats, st = tr.step(T.SceneBatch(gt[v:v + 1], [cams[v]], 1.0), splats)
    cur["e"] = E(); cur["e"].record()
    log.append(cur)
tr.finish()
torch.cuda.synchronize()
from brush_b200 import get_context
ctx = get_context(dev)
ctx.set_timing(True)
for s in range(steps, steps + 3):
    v = view_for_rank(s, world, rank, bench.N_VIEWS)
    splats, st = tr.step(T.SceneBatch(gt[v:v + 1], [cams[v]], 1.0), splats)
    tm = ctx.stage_times_ms(torch.cuda.current_stream().cuda_stream)
ctx.set_timing(False)
if rank == 0:
    print("step stages: overlap=%s " % tr.overlap_exchange + " ".join(f"{k[:12]}={v:.4f}" for k, v in tm.items()) + f" sum_fwd={sum(list(tm.values())[:8]):.4f}", flush=True)
if rank == 0:
    base = log[8]["s"]
    for i in range(8, 13):
        c = log[i]
        row = {k: round(base.elapsed_time(ev), 3) for k, ev in c.items()}
        print(f"step {i} overlap={tr.overlap_exchange}: " + " ".join(f"{k}={v}" for k, v in row.items()), flush=True)
import time
T.render_forward = orig_fwd
tr._peer_step = orig_peer
torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
t0 = time.perf_counter()
for s in range(40):
    v = view_for_rank(s, world, rank, bench.N_VIEWS)
    splats, st = tr.step(T.SceneBatch(gt[v:v + 1], [cams[v]], 1.0), splats)
t1 = time.perf_counter()
tr.finish(); torch.cuda.synchronize()
t2 = time.perf_counter()
if rank == 0:
    print(f"step host: overlap={tr.overlap_exchange} host enqueue {1e3 * (t1 - t0) / 40:.3f} ms/step, total {1e3 * (t2 - t0) / 40:.3f} ms/step", flush=True)
if rank == 0 and os.environ.get("TRACE_PROFILE"):
    import cProfile, pstats
    pr = cProfile.Profile(); pr.enable()
for s in range(40):
    v = view_for_rank(s, world, rank, bench.N_VIEWS)
    splats, st = tr.step(T.SceneBatch(gt[v:v + 1], [cams[v]], 1.0), splats)
if rank == 0 and os.environ.get("TRACE_PROFILE"):
    pr.disable(); torch.cuda.synchronize()
    pstats.Stats(pr).sort_stats("tottime").print_stats(25)
tr.finish(); torch.cuda.synchronize()
dist.destroy_process_group()
