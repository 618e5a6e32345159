"""Bit-exact parity against the reference CUDA build (oracle/_ref) under GENERAL cameras and parameters.

The fixed synthetic configs use a camera at the origin (identity or a small yaw).  The reference's arithmetic is
reproduced operation for operation (fused multiply-adds included), and which products ptxas fuses does not depend on
the data -- but a dense view matrix, an off-centre principal point, a far-away camera, un-normalised quaternions, a
scale modifier and a coloured background exercise every term of every expression with non-trivial values, which the
axis-aligned cameras do not.  Same checks as test_gpu_parity.test_cuda_matches_reference_cuda_bit_exact."""
import math

import pytest
import torch

import htgs_synthetic as syn
from helpers import GRAD_RTOL, bits_equal, rel_inf, run_ours
from oracle import refcuda

pytestmark = [pytest.mark.gpu,
              pytest.mark.skipif(not refcuda.available(), reason="oracle/_ref (reference CUDA build) did not travel to this box")]


def _rot(ax, ay, az):
    cx, sx, cy, sy, cz, sz = math.cos(ax), math.sin(ax), math.cos(ay), math.sin(ay), math.cos(az), math.sin(az)
    Rx = torch.tensor([[1, 0, 0], [0, cx, -sx], [0, sx, cx]], dtype=torch.float64)
    Ry = torch.tensor([[cy, 0, sy], [0, 1, 0], [-sy, 0, cy]], dtype=torch.float64)
    Rz = torch.tensor([[cz, -sz, 0], [sz, cz, 0], [0, 0, 1]], dtype=torch.float64)
    return Rz @ Ry @ Rx


def _general_camera(W, H, seed, device):
    """World->camera = random rotation (up to ~25 deg per axis) about the scene centre plus a random shift; pinhole
    projection with principal point off the image centre (the getProjectionMatrix form of utils/graphics_utils.py)."""
    g = torch.Generator().manual_seed(1000 + seed)
    u = lambda lo, hi: float(lo + (hi - lo) * torch.rand(1, generator=g))
    fovx = u(0.9, 1.6)
    tanfovx = math.tan(0.5 * fovx)
    focal = W / (2.0 * tanfovx)
    tanfovy = H / (2.0 * focal)
    R = _rot(u(-0.45, 0.45), u(-0.45, 0.45), u(-0.45, 0.45))
    pivot = torch.tensor([0.0, 0.0, 4.5], dtype=torch.float64)
    t = pivot - R @ pivot + torch.tensor([u(-0.6, 0.6), u(-0.4, 0.4), u(-0.5, 1.5)], dtype=torch.float64)
    w2c = torch.eye(4, dtype=torch.float64)
    w2c[:3, :3], w2c[:3, 3] = R, t
    view = w2c.t().contiguous().float()
    znear, zfar = 0.01, 100.0
    cx, cy = u(-0.15, 0.15), u(-0.15, 0.15)  # principal point offset in NDC
    proj = torch.zeros(4, 4, dtype=torch.float64)
    proj[0, 0], proj[1, 1] = 1.0 / tanfovx, 1.0 / tanfovy
    proj[0, 2], proj[1, 2] = cx, cy
    proj[3, 2] = 1.0
    proj[2, 2] = zfar / (zfar - znear)
    proj[2, 3] = -(zfar * znear) / (zfar - znear)
    full = (view.unsqueeze(0).bmm(proj.t().contiguous().float().unsqueeze(0))).squeeze(0).contiguous()
    campos = view.inverse()[3, :3].contiguous()
    return dict(W=W, H=H, tanfovx=tanfovx, tanfovy=tanfovy, viewmatrix=view.to(device), projmatrix=full.to(device),
                campos=campos.to(device), focal=focal)


@pytest.mark.parametrize("seed,deg,mod", [(0, 3, 1.0), (1, 3, 0.7), (2, 2, 1.3), (3, 1, 1.0), (4, 0, 1.0), (5, 3, 1.0)])
def test_general_camera_bit_exact_against_reference_cuda(seed, deg, mod):
    dev = torch.device("cuda", 0)
    P, W, H = 30_000, 333, 201  # neither dimension a multiple of the 16-pixel tile
    scene = syn.make_scene(P, W, H, s_max=10.0, seed=40 + seed, device=dev)
    g = torch.Generator().manual_seed(seed)
    # what the rasterizer does NOT normalise must not be assumed normalised: stretch the quaternions
    scene["rotations"] = (scene["rotations"] * (0.5 + torch.rand(P, 1, generator=g).to(dev))).contiguous()
    cam = _general_camera(W, H, seed, dev)
    up = syn.make_upstream(W, H, seed=seed, device=dev)
    bg = torch.tensor([0.25, 0.5, 0.75], device=dev)
    kw = dict(shs=scene["shs"], colors_precomp=None, scales=scene["scales"], rotations=scene["rotations"], cov3D_precomp=None,
              scale_modifier=mod)
    mine = run_ours(scene, cam, up, bg, deg, scale_modifier=mod)
    r = refcuda.forward(scene["means3D"], scene["opacities"], cam, bg, deg, **kw)
    rg = refcuda.backward(r, scene["means3D"], cam, bg, up["dL_dcolor"], up["dL_ddepth"], up["dL_dalpha"], deg, **kw)
    torch.cuda.synchronize()
    vis = r["radii"] > 0
    assert int(vis.sum()) > P // 10, "camera sees too little of the scene for a meaningful comparison"
    assert torch.equal(mine["radii"], r["radii"])
    assert torch.equal(mine["tiles_touched"], r["tiles_touched"])
    assert mine["R"] == r["R"]
    assert torch.equal(mine["keys"], r["point_list_keys"])
    assert torch.equal(mine["point_list"], r["point_list"])
    tiles = mine["ranges"].shape[0]
    assert torch.equal(mine["ranges"], r["ranges"].view(-1, 2)[:tiles])
    assert torch.equal(mine["n_contrib"], r["n_contrib"].view(H, W))
    assert bits_equal(mine["means2D"][vis], r["means2D"].view(P, 2)[vis])
    assert bits_equal(mine["conic_opacity"][vis], r["conic_opacity"].view(P, 4)[vis])
    assert bits_equal(mine["feat"][vis, 3], r["depths"][vis])
    assert bits_equal(mine["feat"][vis, :3], r["rgb"].view(P, 3)[vis])
    for k in ("color", "depth", "alpha"):
        assert bits_equal(mine[k], r[k]), k
    for k in ("dL_dmeans2D", "dL_dopacity", "dL_dmeans3D", "dL_dsh", "dL_dscales", "dL_drotations"):
        assert rel_inf(mine["grads"][k], rg[k].view(mine["grads"][k].shape)) <= GRAD_RTOL, k
