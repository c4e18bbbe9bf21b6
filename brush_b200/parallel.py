"""View-parallel data parallelism for the splat path.  The reference is single-device
(batch size hard-asserted to 1, crates/brush-train/src/train.rs:216-219); this is the layout BASELINE.json
prescribes: every rank holds a full replica of the Gaussians, renders different camera views, and the dense
per-Gaussian gradients are summed before the Adam step.

Four exchange modes:
  * "views": nothing dense is exchanged.  Each rank publishes the nine per-splat sums rasterize_backwards leaves for
             its view (48-byte rows) in peer-mapped memory; every rank redoes the cheap per-splat backward for every
             view and applies Adam to its own full replica (csrc/backward.cu views_adam_kernel).  Cross-GPU traffic
             48*N*(world-1) bytes per rank and step instead of ~2*244*N, one flag barrier.
  * "nccl":  ONE all-reduce over a flat f32 buffer, then the replicated brush_adam_step (the baseline);
  * "hybrid": the SH group (81 % of the dense gradient bytes at SH degree 3) is exchanged as the colour rows it is made
             of (v_r, v_g, v_b, visible: 16 B per splat and view) and rebuilt + Adam-stepped on every rank
             (csrc/backward.cu sh_views_adam_kernel); the 11 remaining floats per splat go through the "peer" kernel.
             Inbound bytes per splat: 16 (world - 1) + 44 -- the smallest of the modes from 3 ranks up.
  * "peer":  ONE brush_dp_adam_step launch per rank over NVLink peer memory: rank r sums its 1/world slice of
             every rank's gradient buffer, applies Adam there (moments sharded), and stores the new parameters into
             every rank's replica (csrc/dp_adam.cu).  PyTorch symmetric memory is only the plumbing that maps the
             buffers into the peers.

The layout / sharding arithmetic is pure host logic, covered by world_size-2 gloo tests on CPU."""
from __future__ import annotations

from typing import List, Sequence, Tuple

import torch


def view_for_rank(step: int, world: int, rank: int, n_views: int) -> int:
    """Index of the training view rank `rank` renders at global step `step`: consecutive views go to consecutive
    ranks, so one step covers `world` distinct views and `n_views / world` steps cover the data set."""
    return (step * world + rank) % n_views


def _pad4(n: int) -> int:
    return (n + 3) // 4 * 4


class FlatLayout:
    """Named segments inside one contiguous f32 buffer; every segment starts on a 16-byte boundary."""

    def __init__(self, names: Sequence[str], shapes: Sequence[Tuple[int, ...]]):
        self.names, self.shapes = list(names), [tuple(s) for s in shapes]
        self.sizes = [int(torch.Size(s).numel()) for s in self.shapes]
        self.offsets, off = [], 0
        for sz in self.sizes:
            self.offsets.append(off)
            off += _pad4(sz)
        self.total = off

    def alloc(self, device) -> torch.Tensor:
        return torch.zeros(self.total, dtype=torch.float32, device=device)

    def views(self, flat: torch.Tensor) -> List[torch.Tensor]:
        assert flat.numel() >= self.total and flat.is_contiguous()
        return [flat.narrow(0, o, s).view(shp) for o, s, shp in zip(self.offsets, self.sizes, self.shapes)]

    def offset(self, name: str) -> int:
        return self.offsets[self.names.index(name)]

    def padded_size(self, name: str) -> int:
        return _pad4(self.sizes[self.names.index(name)])

    def bytes(self) -> int:
        return 4 * self.total


class FlatGradLayout(FlatLayout):
    """The reference's gradient registration order (render.rs:598-624): means 3N | xy 2N | log_scales 3N | quats 4N
    | sh 3KN | opacity N, then the two per-view refine statistics (sum of |v_xy| norms, visible count:
    train.rs:284-316) N | N, then one slot for the rank's loss term.  (13+3K+2)*N floats: 63 N for SH degree 3."""

    NAMES = ("means", "xy", "log_scales", "quats", "sh", "opacity", "grad_norm", "visible", "loss")

    def __init__(self, n: int, k: int):
        self.n, self.k = n, k
        super().__init__(self.NAMES, [(n, 3), (n, 2), (n, 3), (n, 4), (n, k, 3), (n,), (n,), (n,), (4,)])

    def bytes_per_allreduce(self) -> int:
        return 4 * self.total


class FlatParamLayout(FlatLayout):
    """Parameters in the order of the reference's optimizer steps (train.rs:320-356) plus the reduced refine
    statistics the peer exchange hands back: means | opacity | sh | rotation | log_scales | grad_norm | visible."""

    NAMES = ("means", "opacity", "sh", "rotation", "log_scales", "grad_norm", "visible")
    # parameter -> gradient segment
    GRAD_OF = {"means": "means", "opacity": "opacity", "sh": "sh", "rotation": "quats", "log_scales": "log_scales",
               "grad_norm": "grad_norm", "visible": "visible"}

    def __init__(self, n: int, k: int):
        self.n, self.k = n, k
        super().__init__(self.NAMES, [(n, 3), (n,), (n, k, 3), (n, 4), (n, 3), (n,), (n,)])


def owned_range(count: int, world: int, rank: int) -> Tuple[int, int]:
    """The slice of a `count`-float segment rank `rank` owns in the peer exchange: float4-granular, contiguous,
    a partition of [0, count rounded down to 4) over the ranks.  Same arithmetic as brush_dp_owned_range."""
    n4 = count // 4
    return 4 * (n4 * rank // world), 4 * (n4 * (rank + 1) // world)


def allreduce_flat(flat: torch.Tensor, group=None) -> None:
    """Sum over ranks, in place.  NCCL over NVLink/NVSwitch on the GPU box, gloo in the CPU tests."""
    if group is None or torch.distributed.get_world_size(group) == 1:
        return
    torch.distributed.all_reduce(flat, op=torch.distributed.ReduceOp.SUM, group=group)


def global_grad_scale(views_per_rank: int, world: int) -> float:
    """The loss is a mean over the global batch (train.rs:255), so summed gradients are scaled by 1/(B*world)."""
    return 1.0 / float(views_per_rank * world)


class PeerBuffers:
    """The peer-mapped buffers of the "peer" exchange mode: this rank's flat gradient buffer, its flat parameter
    replica and its flag block, each mapped into every rank of `group` through PyTorch symmetric memory, plus the
    raw pointers brush_dp_adam_step needs.  Raises if the group's GPUs cannot map each other's memory."""

    def __init__(self, grad_layout: FlatGradLayout, param_layout: FlatParamLayout, group, device):
        import torch.distributed._symmetric_memory as symm

        self.world, self.rank = group.size(), group.rank()
        if self.world > 8:
            raise ValueError("peer exchange supports at most 8 ranks (one NVSwitch domain)")
        self.grad = symm.empty(grad_layout.total, dtype=torch.float32, device=device)
        self.param = symm.empty(param_layout.total, dtype=torch.float32, device=device)
        self.flags = symm.empty(64, dtype=torch.int32, device=device)
        self.grad.zero_()
        self.param.zero_()
        self.flags.zero_()
        name = group.group_name
        hg, hp, hf = (symm.rendezvous(t, name) for t in (self.grad, self.param, self.flags))
        self._handles = (hg, hp, hf)  # keep the mappings alive
        self.grad_ptrs = [int(p) for p in hg.buffer_ptrs]
        self.param_ptrs = [int(p) for p in hp.buffer_ptrs]
        self.flag_ptrs = [int(p) for p in hf.buffer_ptrs]
        # NVSwitch multicast (multimem.ld_reduce / multimem.st) cuts the per-GPU link traffic by ~world/2 but each
        # operation is slower than a plain peer access: measured on 8xB200 it wins at world 8 (542 vs 691 us for the
        # 1M-splat exchange) and loses at world 2 (627 vs 390 us).  BRUSH_DP_MULTIMEM=0/1 overrides.
        import os

        want = os.environ.get("BRUSH_DP_MULTIMEM")
        use_mc = (self.world >= 4) if want is None else want == "1"
        mc_ok = bool(use_mc and hg.multicast_ptr and hp.multicast_ptr)
        self.grad_mc = int(hg.multicast_ptr) if mc_ok else 0
        self.param_mc = int(hp.multicast_ptr) if mc_ok else 0
        torch.cuda.synchronize(device)
        torch.distributed.barrier(group)  # every rank's zero-fill has landed before anyone's first exchange


VP_STRIDE = 12      # floats per row (csrc/internal.h)
VIEW_TRAILER = 32   # floats behind the rows: the view's RenderUniforms + loss term


class ViewBuffers:
    """The peer-mapped buffers of the "views" exchange mode: two row blocks (alternating by step parity, which is
    what lets one flag barrier per step suffice) and one flag block per rank.  group=None: plain device memory,
    world 1."""

    def __init__(self, n: int, group, device, stride: int = VP_STRIDE):
        """stride = floats per splat row: 12 for the `views` exchange (the nine sums of rasterize_backwards + the
        visibility flag), 4 for the colour rows of the `hybrid` exchange (v_r, v_g, v_b, visible)."""
        from . import _lib

        self.n = n
        self.block = n * stride + VIEW_TRAILER  # floats; 16-byte multiple
        self.world = group.size() if group is not None else 1
        self.rank = group.rank() if group is not None else 0
        if self.world > _lib.DP_MAX_WORLD:
            raise ValueError("views exchange supports at most 8 ranks (one NVSwitch domain)")
        if self.world > 1:
            import torch.distributed._symmetric_memory as symm

            self.rows = symm.empty(2 * self.block, dtype=torch.float32, device=device)
            self.flags = symm.empty(64, dtype=torch.int32, device=device)
            self.rows.zero_()
            self.flags.zero_()
            hr, hf = symm.rendezvous(self.rows, group.group_name), symm.rendezvous(self.flags, group.group_name)
            self._handles = (hr, hf)
            self.row_ptrs = [int(p) for p in hr.buffer_ptrs]
            self.flag_ptrs = [int(p) for p in hf.buffer_ptrs]
            torch.cuda.synchronize(device)
            torch.distributed.barrier(group)
        else:
            self.rows = torch.zeros(2 * self.block, dtype=torch.float32, device=device)
            self.flags = torch.zeros(64, dtype=torch.int32, device=device)
            self.row_ptrs, self.flag_ptrs = [self.rows.data_ptr()], [self.flags.data_ptr()]

    def exchange(self, epoch: int):
        """The BrushViewExchange of step `epoch` (>= 1, the same on every rank)."""
        from . import _lib

        x = _lib.BrushViewExchange()
        x.world, x.rank, x.epoch = self.world, self.rank, epoch
        off = 4 * self.block * (epoch & 1)
        for r in range(self.world):
            x.rows[r], x.flags[r] = self.row_ptrs[r] + off, self.flag_ptrs[r]
        return x
