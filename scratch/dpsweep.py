"""torchrun microbenchmark: brush_dp_adam_step alone, swept over CTAs per SM (BRUSH_DP_U = float4s in flight per thread is
per process).  Prints ms for the whole exchange (all segments) and for the SH segment alone."""
import os, sys, ctypes as C
sys.path.insert(0, '.')
import torch, torch.distributed as dist
rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(rank); dev = torch.device("cuda", rank)
dist.init_process_group("nccl", device_id=dev)
from brush_b200 import Splats, _lib, get_context
from brush_b200.train import SplatTrainer, TrainConfig
N, K = 1_000_000, 16
ctx = get_context(dev)
g = torch.Generator(device=dev).manual_seed(1)
r = lambda *s: torch.randn(*s, device=dev, generator=g)
tr = SplatTrainer(N, TrainConfig(max_refine_step=0), dev, sh_coeffs_per_channel=K, process_group=dist.group.WORLD, exchange="peer")
sp = tr.adopt(Splats(r(N, 3), r(N, K, 3), r(N, 4), r(N), r(N, 3)))
tr.flat_grad.normal_(0, 1e-3)
cfg, pl, gl, peer = tr.config, tr.param_layout, tr.layout, tr.peer
lrs = {"means": 1e-4, "opacity": cfg.lr_opac, "sh": cfg.lr_coeffs_dc, "rotation": cfg.lr_rotation, "log_scales": cfg.lr_scale}
st = torch.cuda.current_stream(dev).cuda_stream
def launch(names, ctas):
    tr.epoch += 1; tr.opt_step += 1
    segs = (_lib.BrushDpSegment * len(names))()
    for i, name in enumerate(names):
        segs[i].src_offset, segs[i].dst_offset, segs[i].count = gl.offset(pl.GRAD_OF[name]), pl.offset(name), pl.padded_size(name)
        segs[i].adam, segs[i].lr = int(name in lrs), lrs.get(name, 0.0)
        segs[i].sh_coeffs_per_channel, segs[i].sh_rest_lr_scale = (K if name == "sh" else 0), 0.05
    s = _lib.BrushDpStep()
    s.world, s.rank = peer.world, peer.rank
    for q in range(peer.world):
        s.src[q], s.dst[q], s.flags[q] = peer.grad_ptrs[q], peer.param_ptrs[q], peer.flag_ptrs[q]
    s.src_multicast, s.dst_multicast = peer.grad_mc or None, peer.param_mc or None
    s.m, s.v = tr.peer_m.data_ptr(), tr.peer_v.data_ptr()
    s.loss_offset, s.loss_out = gl.offset("loss"), tr.loss_buf.data_ptr()
    s.epoch, s.lane, s.ctas_per_sm = tr.epoch, 0, ctas
    ctx.check(_lib.lib().brush_dp_adam_step(ctx.handle, C.byref(s), segs, len(segs), 0.9, 0.999, 1e-15, tr.opt_step, 1.0 / world, C.c_void_p(st)))
def timeit(fn, reps=20):
    for _ in range(3): fn()
    torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps): fn()
    e1.record(); torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) / reps], device=dev); dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t)
allnames = list(pl.NAMES)
for ctas in (1, 2, 3, 4, 6):
    a = timeit(lambda: launch(allnames, ctas))
    b = timeit(lambda: launch(["sh"], ctas))
    if rank == 0:
        print(f"world={world} U={os.environ.get('BRUSH_DP_U','2')} mc={bool(peer.grad_mc)} ctas/SM={ctas}: all {a*1e3:.0f} us, sh only {b*1e3:.0f} us", flush=True)
dist.destroy_process_group()
