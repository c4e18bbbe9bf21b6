"""torchrun microbenchmark of the exchange alone: brush_dp_adam_step (multimem / peer loads) vs NCCL all-reduce + Adam."""
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
st = C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
def mk():
    g = torch.Generator(device=dev).manual_seed(1)
    r = lambda *s: torch.randn(*s, device=dev, generator=g)
    return Splats(r(N, 3), r(N, K, 3), r(N, 4), r(N), r(N, 3))
def timeit(fn, reps=30):
    for _ in range(5): fn()
    torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps): fn()
    e1.record(); torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) / reps], device=dev); dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t)
for mode, force_p2p in (("peer", "mm/mm"), ("peer", "mm/p2p"), ("peer", "p2p/mm"), ("peer", "p2p/p2p"), ("nccl", False)):
    tr = SplatTrainer(N, TrainConfig(max_refine_step=0), dev, sh_coeffs_per_channel=K, process_group=dist.group.WORLD, exchange=mode)
    sp = tr.adopt(mk())
    tr.flat_grad.normal_(0, 1e-3)
    if mode == "peer":
        if force_p2p.startswith("p2p"): tr.peer.grad_mc = 0
        if force_p2p.endswith("p2p"): tr.peer.param_mc = 0
        def fn():
            tr.iter += 1
            tr._peer_step(ctx, 1e-4, 1.0 / world, st)
        name = "peer ld/st=" + force_p2p
    else:
        from brush_b200.parallel import allreduce_flat
        def fn():
            tr.iter += 1
            allreduce_flat(tr.flat_grad, tr.pg)
            tr._adam(ctx, sp, 1e-4, 1.0 / world, st)
        name = "nccl all-reduce + adam"
    ms = timeit(fn)
    if rank == 0: print(f"world={world} {name}: {ms*1e3:.0f} us", flush=True)
    del tr, sp
dist.destroy_process_group()
