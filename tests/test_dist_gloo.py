"""world_size-2 gloo test (CPU) of the view-parallel host logic: view sharding and the single flat
gradient all-reduce used by bench.py --gpus N.  The rasterizer itself is per-rank and not involved."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import htgs_dist


def test_shard_views_partitions():
    for n in (1, 7, 8, 9, 64):
        for world in (1, 2, 4, 8):
            seen = []
            for r in range(world):
                seen += list(htgs_dist.shard_views(n, r, world))
            assert seen == list(range(n))
            sizes = [len(htgs_dist.shard_views(n, r, world)) for r in range(world)]
            assert max(sizes) - min(sizes) <= 1


def _worker(rank, world, port, P, out):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        g = torch.Generator().manual_seed(100 + rank)
        like = dict(means3D=torch.zeros(P, 3), shs=torch.zeros(P, 16, 3), opacities=torch.zeros(P, 1), scales=torch.zeros(P, 3),
                    rotations=torch.zeros(P, 4))
        grads = {k: torch.randn(v.shape, generator=g) for k, v in like.items()}
        grads["scales"] = None if rank == 1 else grads["scales"]  # a rank whose views never touched a group
        red = htgs_dist.FlatGradReducer(P, "cpu")
        red.pack(grads)
        red.all_reduce()
        got = red.unpack(like)
        # expected: recompute both ranks' gradients locally
        exp = {k: torch.zeros_like(v) for k, v in like.items()}
        for r in range(world):
            gr = torch.Generator().manual_seed(100 + r)
            for k, v in like.items():
                t = torch.randn(v.shape, generator=gr)
                if not (k == "scales" and r == 1):
                    exp[k] += t
        ok = all(torch.allclose(got[k], exp[k], atol=1e-6) for k in like)
        out.put((rank, ok, red.flat.numel()))
    finally:
        dist.destroy_process_group()


def test_flat_gradient_all_reduce_world2():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    ctx = mp.get_context("spawn")
    out = ctx.Queue()
    P, world = 257, 2
    procs = [ctx.Process(target=_worker, args=(r, world, port, P, out)) for r in range(world)]
    for p in procs:
        p.start()
    res = [out.get(timeout=120) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert sorted(r for r, _, _ in res) == [0, 1]
    assert all(ok for _, ok, _ in res)
    assert all(n == P * htgs_dist.FLOATS_PER_GAUSSIAN for _, _, n in res)


def _view_outputs(rank, P):
    g = torch.Generator().manual_seed(500 + rank)
    radii = torch.randint(-1, 40, (P,), generator=g, dtype=torch.int32).clamp_(min=0)  # ~5% culled
    radii[torch.rand(P, generator=g) < 0.3] = 0
    grad = torch.randn(P, 3, generator=g)
    return radii, grad


def _stats_worker(rank, world, port, P, out):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        g0 = torch.Generator().manual_seed(9)
        max_r, accum, denom = torch.rand(P, generator=g0) * 20, torch.rand(P, 1, generator=g0), torch.rand(P, 1, generator=g0).round()
        exp_r, exp_a, exp_d = max_r.clone(), accum.clone(), denom.clone()
        radii, grad = _view_outputs(rank, P)
        htgs_dist.reduce_densification_stats(radii, grad, max_r, accum, denom)
        # expected: one process visiting the views of all ranks in turn (the reference's per-view update)
        for r in range(world):
            rr, gg = _view_outputs(r, P)
            vis = rr > 0
            exp_r[vis] = torch.max(exp_r[vis], rr[vis].float())
            exp_a[vis] += torch.norm(gg[vis, :2], dim=-1, keepdim=True)
            exp_d[vis] += 1
        ok = torch.equal(max_r, exp_r) and torch.allclose(accum, exp_a, atol=1e-6) and torch.equal(denom, exp_d)
        out.put((rank, ok, float(accum.sum())))
    finally:
        dist.destroy_process_group()


def test_densification_stats_reduce_world2():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    ctx = mp.get_context("spawn")
    out = ctx.Queue()
    P, world = 1031, 2
    procs = [ctx.Process(target=_stats_worker, args=(r, world, port, P, out)) for r in range(world)]
    for p in procs:
        p.start()
    res = [out.get(timeout=120) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert all(ok for _, ok, _ in res)
    assert res[0][2] == res[1][2]  # replicas identical
