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
    assert lay.total == (13 + 3 * k + 2) * n
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
    assert torch.all(flat == 3.0) and lay.bytes_per_allreduce() == 4 * 18 * 5
