"""Compact SH exchange of view-parallel training (htgs_dist.CompactSHGradReducer, htgs_backward_compact, htgs_sh_expand):
ranks trade the clamp-masked colour gradient (3 floats per Gaussian per view) and rebuild dL_dsh = sum over views of
basis (x) r themselves.  Checked against the ordinary path: the sum of the full dL_dsh tensors of the same views."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import htgs_synthetic as syn
from helpers import GRAD_RTOL, elem_excess, rel_inf

pytestmark = pytest.mark.gpu
NAMES = ("means3D", "shs", "opacities", "scales", "rotations")


def _render_backward(params, cam, up, bg, deg):
    from diff_gaussian_rasterization import GaussianRasterizationSettings, GaussianRasterizer
    rs = GaussianRasterizationSettings(image_height=cam["H"], image_width=cam["W"], tanfovx=cam["tanfovx"], tanfovy=cam["tanfovy"], bg=bg,
                                       scale_modifier=1.0, viewmatrix=cam["viewmatrix"], projmatrix=cam["projmatrix"], sh_degree=deg,
                                       campos=cam["campos"], prefiltered=False, debug=False)
    out = GaussianRasterizer(rs)(means3D=params["means3D"], means2D=torch.zeros_like(params["means3D"], requires_grad=True),
                                 opacities=params["opacities"], shs=params["shs"], scales=params["scales"], rotations=params["rotations"])
    torch.autograd.backward((out[0], out[2], out[3]), (up["dL_dcolor"], up["dL_ddepth"], up["dL_dalpha"]))


@pytest.mark.parametrize("deg", [0, 1, 2, 3])
@pytest.mark.parametrize("cfgname,P", [("small", None), ("ragged", 2999)])  # 2999: P not a multiple of 4 (padded slot rows)
def test_single_rank_two_views_match_ordinary_sum(cfgname, P, deg):
    import htgs_dist
    dev = torch.device("cuda:0")
    cfg, scene, _, up = syn.make_config(cfgname, device=dev)
    if P is not None:
        scene = {k: v[:P].contiguous() for k, v in scene.items()}
    P = scene["means3D"].shape[0]
    W, H = cfg["W"], cfg["H"]
    bg = torch.tensor([0.1, 0.3, 0.2], device=dev)
    cams = [syn.make_camera(W, H, yaw_deg=y, device=dev) for y in (-4.0, 5.0)]
    # ordinary path: autograd accumulates the two views
    plain = {k: v.clone().requires_grad_(True) for k, v in scene.items()}
    for cam in cams:
        _render_backward(plain, cam, up, bg, deg)
    # compact path
    red = htgs_dist.CompactSHGradReducer(P, dev, views_total=2)
    par = {k: v.clone().requires_grad_(True) for k, v in scene.items()}
    red.install(par)
    try:
        for step in range(2):  # twice: slots and arena are reused step after step
            for p in par.values():
                p.grad = None
            red.begin_step()
            for cam in cams:
                _render_backward(par, cam, up, bg, deg)
            assert par["shs"].grad is None  # the expanded SH gradient was never produced by the backward
            assert red.pack_if_needed({k: par[k].grad for k in htgs_dist.SMALL_ORDER}) == 0
            red.all_reduce(par["means3D"], deg)
            got = red.grads(par)
            torch.cuda.synchronize()
            for k in NAMES:
                assert rel_inf(got[k], plain[k].grad) <= GRAD_RTOL, (k, step)
                assert elem_excess(got[k], plain[k].grad) <= 1.0, (k, step)
            act = (deg + 1) ** 2
            assert not got["shs"][:, act:].any()  # coefficients above the active degree stay zero
        with pytest.raises(RuntimeError):  # a third backward without begin_step() has no slot to write to
            _render_backward(par, cams[0], up, bg, deg)
    finally:
        red.remove()


def _worker(rank, world, port, mode, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world), LOCAL_RANK=str(rank))
    import importlib
    importlib.import_module("3dgs_hierarchical_training_b200").install()
    import htgs_dist
    dev = torch.device("cuda", rank)
    torch.cuda.set_device(dev)
    dist.init_process_group("nccl", device_id=dev)
    try:
        cfg, scene, _, up = syn.make_config("small", device=dev)
        P, W, H = cfg["P"], cfg["W"], cfg["H"]
        bg = torch.zeros(3, device=dev)
        views = 2 * world  # two views per rank
        cams = [syn.make_camera(W, H, yaw_deg=2.0 * v - views, device=dev) for v in range(views)]
        mine = cams[2 * rank:2 * rank + 2]
        # reference: ordinary gradients of the local views, NCCL all-reduce of the flat 59-float buffer
        plain = {k: v.clone().requires_grad_(True) for k, v in scene.items()}
        for cam in mine:
            _render_backward(plain, cam, up, bg, 3)
        ref = htgs_dist.FlatGradReducer(P, dev)
        ref.pack({k: plain[k].grad for k in htgs_dist.PARAM_ORDER})
        ref.all_reduce()
        want = ref.unpack(plain)
        red = htgs_dist.CompactSHGradReducer(P, dev, views_total=views, peer_memory=mode)
        par = {k: v.clone().requires_grad_(True) for k, v in scene.items()}
        red.install(par)
        ok = True
        try:
            for step in range(2):
                for p in par.values():
                    p.grad = None
                red.begin_step()
                for cam in mine:
                    _render_backward(par, cam, up, bg, 3)
                red.pack_if_needed({k: par[k].grad for k in htgs_dist.SMALL_ORDER})
                red.all_reduce(par["means3D"], 3)
                got = red.grads(par)
                torch.cuda.synchronize()
                for k in NAMES:
                    ok = ok and rel_inf(got[k], want[k]) <= GRAD_RTOL and elem_excess(got[k], want[k]) <= 1.0
                # replicas must hold the same bits
                lo, hi = red.sh_grad.clone(), red.sh_grad.clone()
                dist.all_reduce(lo, op=dist.ReduceOp.MIN)
                dist.all_reduce(hi, op=dist.ReduceOp.MAX)
                ok = ok and torch.equal(lo, hi)
        finally:
            red.remove()
        out.put((rank, bool(ok), red.collective))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 4])
@pytest.mark.parametrize("mode", [False, "p2p", "multimem"])
def test_multi_rank_matches_flat_allreduce(mode, world):
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    ctx = mp.get_context("spawn")
    out = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, world, port, mode, out)) for r in range(world)]
    for p in procs:
        p.start()
    res = [out.get(timeout=300) for _ in range(world)]
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    assert all(ok for _, ok, _ in res)
    assert all(c == (mode or "nccl") for _, _, c in res)
