"""View-parallel data parallelism for the splat path.  The reference is single-device
(batch size hard-asserted to 1, crates/brush-train/src/train.rs:216-219); this is the layout BASELINE.json
prescribes: every rank holds a full replica of the Gaussians, renders different camera views, and the dense
per-Gaussian gradients are summed with ONE all-reduce over a flat f32 buffer before the replicated Adam step.

Pure-torch host logic (no CUDA needed), so it is covered by world_size-2 gloo tests on CPU."""
from __future__ import annotations

from typing import List, Tuple

import torch


def view_for_rank(step: int, world: int, rank: int, n_views: int) -> int:
    """Index of the training view rank `rank` renders at global step `step`: consecutive views go to consecutive
    ranks, so one step covers `world` distinct views and `n_views / world` steps cover the data set."""
    return (step * world + rank) % n_views


class FlatGradLayout:
    """One contiguous f32 buffer holding, in the reference's gradient registration order (render.rs:598-624),
    means 3N | xy 2N | log_scales 3N | quats 4N | sh 3KN | opacity N, then the two per-view refine statistics
    (sum of |v_xy| norms, visible count: train.rs:284-316) N | N.  (13+3K+2)*N floats: 63 N for SH degree 3."""

    NAMES = ("means", "xy", "log_scales", "quats", "sh", "opacity", "grad_norm", "visible")

    def __init__(self, n: int, k: int):
        self.n, self.k = n, k
        self.sizes = [3 * n, 2 * n, 3 * n, 4 * n, 3 * k * n, n, n, n]
        self.shapes = [(n, 3), (n, 2), (n, 3), (n, 4), (n, k, 3), (n,), (n,), (n,)]
        self.total = sum(self.sizes)

    def alloc(self, device) -> torch.Tensor:
        return torch.zeros(self.total, dtype=torch.float32, device=device)

    def views(self, flat: torch.Tensor) -> List[torch.Tensor]:
        assert flat.numel() == self.total and flat.is_contiguous()
        return [p.view(s) for p, s in zip(torch.split(flat, self.sizes), self.shapes)]

    def bytes_per_allreduce(self) -> int:
        return 4 * self.total


def allreduce_flat(flat: torch.Tensor, group=None) -> None:
    """Sum over ranks, in place.  NCCL over NVLink/NVSwitch on the GPU box, gloo in the CPU tests."""
    if group is None or torch.distributed.get_world_size(group) == 1:
        return
    torch.distributed.all_reduce(flat, op=torch.distributed.ReduceOp.SUM, group=group)


def global_grad_scale(views_per_rank: int, world: int) -> float:
    """The loss is a mean over the global batch (train.rs:255), so summed gradients are scaled by 1/(B*world)."""
    return 1.0 / float(views_per_rank * world)
