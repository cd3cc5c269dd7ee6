"""View-sharded data parallelism for the splat path.  One process per GPU (torchrun).

The reference is single-GPU; the path shards by view: parameters are replicated, each rank renders its own views,
and the Gaussian-parameter gradients are summed with ONE all-reduce over a flat bucket (NCCL over NVLink on the
GPU box, gloo in the CPU tests).  Densification statistics need a sum / max all-reduce of [P] tensors so every rank
takes the same densify decisions (train.py:130,144; scene/gaussian_model.py:627-629).  The relit-eval grid
(view x envmap cells) shards view-major with no collective, so each rank keeps the geometry of its views cached.
"""
from typing import List, Sequence, Tuple

import torch


def views_for_rank(n_views: int, rank: int, world: int) -> List[int]:
    """Training: rank r takes views r, r+world, ... (round robin keeps per-step work balanced)."""
    return list(range(rank, n_views, world))


def grid_cells_for_rank(n_views: int, n_envmaps: int, rank: int, world: int) -> List[Tuple[int, int]]:
    """Eval grid: contiguous VIEW-MAJOR blocks, so a rank loops envmaps inside a view (geometry cached)."""
    per = (n_views + world - 1) // world
    lo, hi = min(rank * per, n_views), min((rank + 1) * per, n_views)
    return [(v, e) for v in range(lo, hi) for e in range(n_envmaps)]


class GradBucket:
    """Flat fp32 bucket over a fixed parameter list: pack -> all_reduce(sum) -> unpack (optionally averaged)."""

    def __init__(self, params: Sequence[torch.Tensor]):
        self.params = list(params)
        self.sizes = [p.numel() for p in self.params]
        self.total = sum(self.sizes)
        self.flat = torch.zeros(self.total, dtype=torch.float32, device=self.params[0].device)

    def pack(self):
        """Gradients -> flat bucket: ONE multi-tensor copy (parameters without a gradient contribute zeros)."""
        views = self.flat.split(self.sizes)
        dst, src = [], []
        for v, p in zip(views, self.params):
            if p.grad is None:
                v.zero_()
            elif p.grad.data_ptr() != v.data_ptr():      # already a bucket view (gradient accumulated in place)
                dst.append(v)
                src.append(p.grad.reshape(-1))
        if dst:
            torch._foreach_copy_(dst, src)
        return self.flat

    def unpack(self, scale: float = 1.0):
        """Bucket -> gradients WITHOUT a copy: every p.grad becomes a view of the flat bucket (as DDP's
        gradient_as_bucket_view), valid until the next pack(); the scale is one in-place kernel over the bucket."""
        if scale != 1.0:
            self.flat.mul_(scale)
        for v, p in zip(self.flat.split(self.sizes), self.params):
            p.grad = v.view_as(p)

    def all_reduce(self, average_over: int = 0):
        """pack -> ONE all-reduce -> unpack.  With NCCL and average_over == world size the division rides inside the
        collective (ReduceOp.AVG): per step 1 multi-tensor copy + 1 collective, no per-parameter kernels."""
        import torch.distributed as dist
        self.pack()
        scale = 1.0 / average_over if average_over else 1.0
        if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
            if average_over == dist.get_world_size() and dist.get_backend() == "nccl":
                dist.all_reduce(self.flat, op=dist.ReduceOp.AVG)
                scale = 1.0
            else:
                dist.all_reduce(self.flat, op=dist.ReduceOp.SUM)
        self.unpack(scale)


def sync_densification_stats(xyz_gradient_accum: torch.Tensor, denom: torch.Tensor, max_radii2D: torch.Tensor):
    """sum / sum / max all-reduce so densify-and-prune is identical on every rank."""
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1):
        return
    dist.all_reduce(xyz_gradient_accum, op=dist.ReduceOp.SUM)
    dist.all_reduce(denom, op=dist.ReduceOp.SUM)
    dist.all_reduce(max_radii2D, op=dist.ReduceOp.MAX)
