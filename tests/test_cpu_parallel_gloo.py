"""World-size-2 gloo tests (CPU) of the view-sharding host logic: partitioning, gradient bucket, densification stats."""
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
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    sys.path.insert(0, os.path.join(root, "relightable3dgaussians-w_b200"))
    from r3dgw_b200 import parallel as par
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        g = torch.Generator().manual_seed(5)
        params = [torch.randn(50, 3, generator=g).requires_grad_(True), torch.randn(50, 1, generator=g).requires_grad_(True),
                  torch.randn(25, 3, generator=g).requires_grad_(True)]
        views = par.views_for_rank(8, rank, world)
        # a per-view "loss" whose gradient is known: sum_v (v+1) * sum(p^2)  -> grad = 2 (v+1) p
        loss = sum((v + 1) * sum((p * p).sum() for p in params[:2]) for v in views)
        loss.backward()                     # params[2] gets no grad on purpose (bucket must send zeros)
        bucket = par.GradBucket(params)
        bucket.all_reduce(average_over=8)
        expect_scale = 2.0 * sum(v + 1 for v in range(8)) / 8.0
        ok = all(torch.allclose(p.grad, expect_scale * p.detach(), rtol=1e-6, atol=1e-6) for p in params[:2])
        ok = ok and bool((params[2].grad == 0).all())
        acc = torch.full((10,), float(rank + 1)); den = torch.ones(10); mx = torch.arange(10.0) * (rank + 1)
        par.sync_densification_stats(acc, den, mx)
        ok = ok and bool((acc == 3).all()) and bool((den == 2).all()) and torch.equal(mx, torch.arange(10.0) * 2)
        cells = par.grid_cells_for_rank(5, 3, rank, world)
        gathered = [None] * world
        dist.all_gather_object(gathered, cells)
        flat = [c for cs in gathered for c in cs]
        ok = ok and sorted(flat) == [(v, e) for v in range(5) for e in range(3)] and len(set(flat)) == 15
        # view-major: a rank's cells of one view are contiguous
        ok = ok and all(cells[i][0] <= cells[i + 1][0] for i in range(len(cells) - 1))
        out[rank] = ok
    finally:
        dist.destroy_process_group()


def test_view_sharding_gloo_world2():
    world = 2
    port = _free_port()
    mgr = mp.Manager()
    out = mgr.dict()
    mp.spawn(_worker, args=(world, port, out), nprocs=world, join=True)
    assert dict(out) == {0: True, 1: True}


def test_partition_helpers_single_process():
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    sys.path.insert(0, os.path.join(root, "relightable3dgaussians-w_b200"))
    from r3dgw_b200 import parallel as par
    for world in (1, 2, 4, 8):
        allv = sorted(v for r in range(world) for v in par.views_for_rank(8, r, world))
        assert allv == list(range(8))
        cells = [c for r in range(world) for c in par.grid_cells_for_rank(64, 8, r, world)]
        assert len(cells) == 512 and len(set(cells)) == 512
    assert par.grid_cells_for_rank(3, 2, 7, 8) == []      # more ranks than views: empty shard, not an error
