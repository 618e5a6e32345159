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


# ---- host logic of the compact SH exchange (slot ownership, all-gather order, small all-reduce) on gloo ----

def _basis(dirs, deg):
    """Real SH basis in the reference's convention (utils/sh_utils.py:57-112 constants), numpy-free torch version."""
    x, y, z = dirs[:, 0], dirs[:, 1], dirs[:, 2]
    C0, C1 = 0.28209479177387814, 0.4886025119029199
    C2 = [1.0925484305920792, -1.0925484305920792, 0.31539156525252005, -1.0925484305920792, 0.5462742152960396]
    C3 = [-0.5900435899266435, 2.890611442640554, -0.4570457994644658, 0.3731763325901154, -0.4570457994644658, 1.445305721320277,
          -0.5900435899266435]
    w = torch.zeros(dirs.shape[0], 16, dtype=torch.float64)
    w[:, 0] = C0
    if deg > 0:
        w[:, 1], w[:, 2], w[:, 3] = -C1 * y, C1 * z, -C1 * x
    if deg > 1:
        xx, yy, zz, xy, yz, xz = x * x, y * y, z * z, x * y, y * z, x * z
        w[:, 4], w[:, 5], w[:, 6], w[:, 7], w[:, 8] = C2[0] * xy, C2[1] * yz, C2[2] * (2 * zz - xx - yy), C2[3] * xz, C2[4] * (xx - yy)
    if deg > 2:
        w[:, 9], w[:, 10], w[:, 11] = C3[0] * y * (3 * xx - yy), C3[1] * xy * z, C3[2] * y * (4 * zz - xx - yy)
        w[:, 12], w[:, 13] = C3[3] * z * (2 * zz - 3 * xx - 3 * yy), C3[4] * x * (4 * zz - xx - yy)
        w[:, 14], w[:, 15] = C3[5] * z * (xx - yy), C3[6] * x * (xx - 3 * yy)
    return w


def _view_rgb(view, P):
    g = torch.Generator().manual_seed(900 + view)
    r = torch.randn(P, 3, generator=g)
    r[torch.rand(P, generator=g) < 0.3] = 0.0  # culled in this view
    return r, torch.randn(3, generator=g) * 5.0  # colour gradient, camera position


def _compact_worker(rank, world, port, P, views, out):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        red = htgs_dist.CompactSHGradReducer(P, "cpu", views_total=views, peer_memory=False)
        means = torch.randn(P, 3, generator=torch.Generator().manual_seed(1))
        P4 = (P + 3) // 4 * 4

        def host_expand(slot_addresses, means3D, deg):  # stands in for the CUDA kernel: same contract, read from the slots
            acc = torch.zeros(P, 16, 3, dtype=torch.float64)
            for v in range(views):
                sl = red.slot(v)
                r, cam = sl[:3 * P].view(P, 3).double(), sl[3 * P4:3 * P4 + 3].double()
                d = means3D.double() - cam
                acc += _basis(d / d.norm(dim=1, keepdim=True), deg)[:, :, None] * r[:, None, :]
            red.sh_grad.copy_(acc.float())

        red._expand = host_expand
        ok = True
        for step in range(2):
            red.begin_step()
            small = {}
            for name, per in zip(htgs_dist.SMALL_ORDER, htgs_dist.SMALL_FLOATS):
                small[name] = torch.randn(P, per, generator=torch.Generator().manual_seed(10 * step + 7 * rank + per))
            for i in range(red.views_per_rank):  # what the backward of each local view leaves behind
                ptr, mc = red._provider("cpu")
                view = rank * red.views_per_rank + i
                assert not mc and ptr == red.slot(view).data_ptr()
                r, cam = _view_rgb(view + 100 * step, P)
                sl = red.slot(view)
                sl[:3 * P] = r.reshape(-1)
                sl[3 * P4:3 * P4 + 3] = cam
            assert red.pack_if_needed(small) == len(htgs_dist.SMALL_ORDER)
            red.all_reduce(means, 3)
            got = red.grads(dict(small, shs=torch.zeros(P, 16, 3)))
            # expected, rebuilt locally from every rank's seeds
            for name, per in zip(htgs_dist.SMALL_ORDER, htgs_dist.SMALL_FLOATS):
                want = sum(torch.randn(P, per, generator=torch.Generator().manual_seed(10 * step + 7 * r_ + per)) for r_ in range(world))
                ok = ok and torch.allclose(got[name], want, atol=1e-6)
            want_sh = torch.zeros(P, 16, 3, dtype=torch.float64)
            for v in range(views):
                r, cam = _view_rgb(v + 100 * step, P)
                d = means.double() - cam.double()
                want_sh += _basis(d / d.norm(dim=1, keepdim=True), 3)[:, :, None] * r.double()[:, None, :]
            ok = ok and torch.allclose(got["shs"].double(), want_sh, atol=1e-5)
        try:
            red._provider("cpu")
            ok = False  # a further backward without begin_step() must be refused
        except RuntimeError:
            pass
        out.put((rank, bool(ok), float(red.sh_grad.double().sum())))
    finally:
        dist.destroy_process_group()


def test_compact_sh_exchange_host_logic_world2():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    ctx = mp.get_context("spawn")
    out = ctx.Queue()
    P, world, views = 1031, 2, 4  # P not a multiple of 4: padded slot rows; two views per rank
    procs = [ctx.Process(target=_compact_worker, args=(r, world, port, P, views, out)) for r in range(world)]
    for p in procs:
        p.start()
    res = [out.get(timeout=180) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert all(ok for _, ok, _ in res)
    assert res[0][2] == res[1][2]  # replicas identical
