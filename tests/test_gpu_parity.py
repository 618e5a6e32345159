"""GPU parity tests proper: the hand-written CUDA path, called through the plugin / C ABI, against
  (1) the CPU oracle (oracle/htgs_oracle.c) on seeded scenes the oracle finishes in seconds,
  (2) the UNMODIFIED reference CUDA rasterizer (oracle/_ref, built from /root/reference) when the
      prebuilt library travelled to the box -- the strict bar: every integer output AND the images
      bit-identical, gradients within 1e-4 relative (the reference's atomics are order-nondeterministic),
  (3) the committed golden vectors produced by (2).
Edge cases follow SURVEY.md Appendix B."""
import glob
import math
import os

import numpy as np
import pytest
import torch

import htgs_synthetic as syn
from helpers import (GRAD_NAMES, GRAD_RTOL, IMAGE_ATOL, ORACLE_GRAD_ELEM_RTOL, ORACLE_GRAD_RTOL, ORACLE_PIXEL_OUTLIER_FRAC, bits_equal, cov3d_of,
                     elem_excess, np_, rel_inf, run_ours)
from oracle import golden, refcuda

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def dev():
    return torch.device("cuda:0")


def _case(cfgname, deg=3, bgv=(0.3, 0.6, 0.9), precomp=False, use_cov=False, mod=1.0):
    cfg, scene, cam, up = syn.make_config(cfgname, device=dev())
    P = cfg["P"]
    bg = torch.tensor(bgv, dtype=torch.float32, device=dev())
    colors = torch.rand(P, 3, generator=torch.Generator().manual_seed(5)).to(dev()) if precomp else None
    cov = cov3d_of({k: v.cpu() for k, v in scene.items()}).to(dev()) if use_cov else None
    kw = dict(shs=None if precomp else scene["shs"], colors_precomp=colors, scales=None if use_cov else scene["scales"],
              rotations=None if use_cov else scene["rotations"], cov3D_precomp=cov, scale_modifier=mod)
    return cfg, scene, cam, up, bg, colors, cov, kw


def _grad_list(precomp, use_cov):
    names = ["dL_dmeans2D", "dL_dopacity", "dL_dmeans3D", "dL_dcov3D"]
    names += ["dL_dcolors"] if precomp else ["dL_dsh"]
    names += [] if use_cov else ["dL_dscales", "dL_drotations"]
    return names


CASES = [("tiny", 3, False, False, 1.0), ("small", 3, False, False, 1.0), ("ragged", 3, False, False, 0.8),
         ("deep", 3, False, False, 1.0), ("small", 0, False, False, 1.0), ("small", 1, False, False, 1.0),
         ("small", 2, False, False, 1.0), ("small", 3, True, False, 1.0), ("small", 3, True, True, 1.0),
         ("config1", 3, False, False, 1.0)]


@pytest.mark.parametrize("cfgname,deg,precomp,use_cov,mod", CASES)
def test_cuda_matches_oracle(cfgname, deg, precomp, use_cov, mod):
    cfg, scene, cam, up, bg, colors, cov, kw = _case(cfgname, deg, precomp=precomp, use_cov=use_cov, mod=mod)
    P, W, H = cfg["P"], cfg["W"], cfg["H"]
    mine = run_ours(scene, cam, up, bg, deg, colors_precomp=colors, cov3D_precomp=cov, scale_modifier=mod)
    o = golden.forward(scene["means3D"], scene["opacities"], cam, bg, deg, **kw)
    og = golden.backward(o, up["dL_dcolor"], up["dL_ddepth"], up["dL_dalpha"])
    vis = o["radii"] > 0
    # bit-exact: integer/index outputs and the fp32 per-Gaussian state they derive from
    assert np.array_equal(np_(mine["radii"]), o["radii"])
    assert np.array_equal(np_(mine["tiles_touched"]).view(np.uint32), o["tiles_touched"])
    assert np.array_equal(np_(mine["point_offsets"]).view(np.uint32), o["point_offsets"])
    assert mine["R"] == o["R"]
    assert np.array_equal(np_(mine["keys"]).view(np.uint64), o["keys"])
    assert np.array_equal(np_(mine["point_list"]).view(np.uint32), o["point_list"])
    assert np.array_equal(np_(mine["ranges"]).view(np.uint32), o["ranges"])
    assert bits_equal(np_(mine["means2D"])[vis], o["means2D"][vis])
    assert bits_equal(np_(mine["conic_opacity"])[vis], o["conic_opacity"][vis])
    assert bits_equal(np_(mine["feat"])[vis, 3], o["depths"][vis])
    if not precomp:
        assert bits_equal(np_(mine["feat"])[vis, :3], o["rgb"][vis])
    # images: 1e-5, with the documented oracle-only allowance for ex2.approx threshold flips
    budget = max(1, int(math.ceil(ORACLE_PIXEL_OUTLIER_FRAC * W * H)))
    nflip = int((np_(mine["n_contrib"]).view(np.uint32) != o["n_contrib"]).sum())
    assert nflip <= budget
    flipped = nflip > 0
    for k in ("color", "depth", "alpha"):
        bad = np.abs(np_(mine[k]) - o[k]) > IMAGE_ATOL * max(1.0, np.abs(o[k]).max())
        assert bad.sum() <= budget * o[k].shape[0], f"{k}: {bad.sum()} values beyond tolerance"
        flipped |= bool(bad.any())
    tol = ORACLE_GRAD_RTOL if flipped else GRAD_RTOL
    for k in _grad_list(precomp, use_cov):
        assert rel_inf(mine["grads"][k], og[k].reshape(np_(mine["grads"][k]).shape)) <= tol, k
        if not flipped:  # element-wise bar too, unless an ex2.approx threshold flip moved a whole contribution
            assert elem_excess(mine["grads"][k], og[k].reshape(np_(mine["grads"][k]).shape), rtol=ORACLE_GRAD_ELEM_RTOL) <= 1.0, k


# the BASELINE.json shapes themselves: config2 (300k, 960x540), config3 (1M, 1080p, the headline), config5 (2M, 512x512
# depth-complexity stress: where the reach masks and the batch skipping are most exposed), config4 (6M, 1080p, one view)
BASELINE_CASES = [("config2", 3, False, False, 1.0), ("config3", 3, False, False, 1.0), ("config5", 3, False, False, 1.0),
                  ("config4", 3, False, False, 1.0)]


@pytest.mark.skipif(not refcuda.available(), reason="oracle/_ref (reference CUDA build) did not travel to this box")
@pytest.mark.parametrize("cfgname,deg,precomp,use_cov,mod", CASES + BASELINE_CASES)
def test_cuda_matches_reference_cuda_bit_exact(cfgname, deg, precomp, use_cov, mod):
    cfg, scene, cam, up, bg, colors, cov, kw = _case(cfgname, deg, precomp=precomp, use_cov=use_cov, mod=mod)
    P, W, H = cfg["P"], cfg["W"], cfg["H"]
    mine = run_ours(scene, cam, up, bg, deg, colors_precomp=colors, cov3D_precomp=cov, scale_modifier=mod)
    r = refcuda.forward(scene["means3D"], scene["opacities"], cam, bg, deg, **kw)
    rg = refcuda.backward(r, scene["means3D"], cam, bg, up["dL_dcolor"], up["dL_ddepth"], up["dL_dalpha"], deg, **kw)
    torch.cuda.synchronize()
    vis = r["radii"] > 0
    assert torch.equal(mine["radii"], r["radii"])
    assert torch.equal(mine["tiles_touched"], r["tiles_touched"])
    assert torch.equal(mine["point_offsets"], r["point_offsets"])
    assert mine["R"] == r["R"]
    assert torch.equal(mine["keys"], r["point_list_keys"])          # sorted 64-bit tile|depth keys
    assert torch.equal(mine["point_list"], r["point_list"])        # sorted Gaussian ids (stability included)
    tiles = mine["ranges"].shape[0]
    assert torch.equal(mine["ranges"], r["ranges"].view(-1, 2)[:tiles])
    assert torch.equal(mine["n_contrib"], r["n_contrib"].view(H, W))
    assert bits_equal(mine["means2D"][vis], r["means2D"].view(P, 2)[vis])
    assert bits_equal(mine["conic_opacity"][vis], r["conic_opacity"].view(P, 4)[vis])
    assert bits_equal(mine["feat"][vis, 3], r["depths"][vis])
    if not precomp:
        assert bits_equal(mine["feat"][vis, :3], r["rgb"].view(P, 3)[vis])
        mybits = ((mine["clamped"][:, None] >> torch.arange(3, device=dev())[None]) & 1).to(torch.uint8)
        assert torch.equal(mybits[vis], r["clamped"].view(P, 3)[vis])
    # images: bit-identical (stricter than the 1e-5 bar)
    assert bits_equal(mine["color"], r["color"])
    assert bits_equal(mine["depth"], r["depth"])
    assert bits_equal(mine["alpha"], r["alpha"])
    for k in _grad_list(precomp, use_cov):
        assert rel_inf(mine["grads"][k], rg[k].view(mine["grads"][k].shape)) <= GRAD_RTOL, k
        assert elem_excess(mine["grads"][k], rg[k].view(mine["grads"][k].shape)) <= 1.0, k
    refcuda.dgr().ref_release()  # give the reference's geometry / binning / image buffers back (6M Gaussians: ~1.5 GB)


# Gradients against the live reference CUDA under other backgrounds than the one above.  Found with tools/grad_margin.py:
# with background (0.1, 0.2, 0.3) at config 5 the first tensor-core form of the backward (moments about the TILE centre,
# shifted to the Gaussian's centre when the batch is flushed) left dL_dcov3D of a sub-pixel Gaussian near a tile corner
# at 1.35-1.7x the element-wise bar while every committed case passed.  Since the shift happens per warp from its 8x4
# block's centre, all tensors sit at <= 0.16 of the bar and <= 3e-6 of each tensor's largest entry on configs 2 / 3 / 5;
# the assertions below hold them to a quarter of the bar and to 1e-5.
@pytest.mark.skipif(not refcuda.available(), reason="oracle/_ref (reference CUDA build) did not travel to this box")
@pytest.mark.parametrize("cfgname,bgv", [("config5", (0.1, 0.2, 0.3)), ("config5", (0.0, 0.0, 0.0)), ("config3", (0.1, 0.2, 0.3)),
                                         ("config2", (1.0, 1.0, 1.0))])
def test_gradients_vs_reference_cuda_other_backgrounds(cfgname, bgv):
    cfg, scene, cam, up, bg, colors, cov, kw = _case(cfgname, 3, bgv=bgv)
    mine = run_ours(scene, cam, up, bg, 3)
    r = refcuda.forward(scene["means3D"], scene["opacities"], cam, bg, 3, **kw)
    rg = refcuda.backward(r, scene["means3D"], cam, bg, up["dL_dcolor"], up["dL_ddepth"], up["dL_dalpha"], 3, **kw)
    torch.cuda.synchronize()
    assert bits_equal(mine["color"], r["color"])
    for k in _grad_list(False, False):
        assert rel_inf(mine["grads"][k], rg[k].view(mine["grads"][k].shape)) <= 1e-5, k
        assert elem_excess(mine["grads"][k], rg[k].view(mine["grads"][k].shape)) <= 0.25, k
    refcuda.dgr().ref_release()


REFCUDA_CASES = [p for p in sorted(glob.glob(os.path.join(GOLD, "refcuda_*.npz"))) if not p.endswith("refcuda_knn.npz")]


@pytest.mark.parametrize("path", REFCUDA_CASES, ids=[os.path.basename(p)[8:-4] for p in REFCUDA_CASES])
def test_cuda_matches_committed_reference_goldens(path):
    v = np.load(path)
    cfgname, deg, precomp, use_cov, mod = v["case"]
    deg, precomp, use_cov, mod = int(deg), precomp == "True", use_cov == "True", float(mod)
    cfg, scene, cam, up, bg, colors, cov, kw = _case(str(cfgname), deg, tuple(v["bg"]), precomp, use_cov, mod)
    P, W, H = cfg["P"], cfg["W"], cfg["H"]
    mine = run_ours(scene, cam, up, bg, deg, colors_precomp=colors, cov3D_precomp=cov, scale_modifier=mod)
    assert np.array_equal(np_(mine["radii"]), v["radii"])
    assert np.array_equal(np_(mine["tiles_touched"]), v["tiles_touched"])
    assert np.array_equal(np_(mine["keys"]), v["point_list_keys"])
    assert np.array_equal(np_(mine["point_list"]), v["point_list"])
    tiles = mine["ranges"].shape[0]
    assert np.array_equal(np_(mine["ranges"]), v["ranges"].reshape(-1, 2)[:tiles])
    assert np.array_equal(np_(mine["n_contrib"]), v["n_contrib"].reshape(H, W))
    for k in ("color", "depth", "alpha"):
        assert bits_equal(mine[k], v[k]), k
    for k in _grad_list(precomp, use_cov):
        assert rel_inf(mine["grads"][k], v[k].reshape(np_(mine["grads"][k]).shape)) <= GRAD_RTOL, k
        assert elem_excess(mine["grads"][k], v[k].reshape(np_(mine["grads"][k]).shape)) <= 1.0, k


def test_full_size_properties_config3():
    """BASELINE headline size (1M Gaussians, 1080p): size-independent properties instead of an oracle run."""
    cfg, scene, cam, up = syn.make_config("config3", device=dev())
    P, W, H = cfg["P"], cfg["W"], cfg["H"]
    bg = torch.zeros(3, device=dev())
    a = run_ours(scene, cam, up, bg)
    k, pl = a["keys"], a["point_list"]
    assert bool((k[1:] >= k[:-1]).all()), "keys not sorted"
    tie = k[1:] == k[:-1]
    assert bool((pl[1:][tie] > pl[:-1][tie]).all()), "sort not stable on (tile, depth) ties"
    assert int(a["point_offsets"][-1]) == a["R"] == int(a["tiles_touched"].sum())
    assert torch.equal(torch.cumsum(a["tiles_touched"], 0).int(), a["point_offsets"])
    # the value multiset survives the sort: every Gaussian appears tiles_touched times
    cnt = torch.bincount(pl.long(), minlength=P)
    assert torch.equal(cnt.int(), a["tiles_touched"])
    r = a["ranges"].long()
    ne = r[:, 1] > r[:, 0]
    assert int((r[ne, 1] - r[ne, 0]).sum()) == a["R"]
    first_tile = (k[r[ne, 0]] >> 32)
    assert torch.equal(first_tile, torch.nonzero(ne).flatten())
    # key = tile<<32 | depth bits of the Gaussian it names
    depth_bits = a["feat"][:, 3].contiguous().view(torch.int32).long() & 0xFFFFFFFF
    assert torch.equal(k & 0xFFFFFFFF, depth_bits[pl.long()])
    # determinism of the forward (idempotence) and finiteness of everything
    b = run_ours(scene, cam, up, bg)
    assert torch.equal(a["color"], b["color"]) and torch.equal(a["point_list"], b["point_list"])
    for t in (a["color"], a["depth"], a["alpha"], *a["grads"].values()):
        assert bool(torch.isfinite(t).all())
    assert float(a["alpha"].max()) <= 1.0 + 1e-5 and float(a["alpha"].min()) >= 0.0
    # linearity of the backward in the upstream gradient: bwd(2u) == 2*bwd(u) up to atomic-order noise
    up2 = {k_: 2.0 * v for k_, v in up.items()}
    c = run_ours(scene, cam, up2, bg)
    for name in ("dL_dmeans3D", "dL_dsh", "dL_dopacity", "dL_dscales", "dL_drotations"):
        assert rel_inf(c["grads"][name], 2.0 * a["grads"][name]) <= GRAD_RTOL, name


def test_edge_cases():
    from diff_gaussian_rasterization import GaussianRasterizationSettings, GaussianRasterizer
    cfg, scene, cam, up = syn.make_config("tiny", device=dev())
    W, H = cfg["W"], cfg["H"]
    bg = torch.tensor([0.2, 0.4, 0.6], device=dev())
    rs = GaussianRasterizationSettings(H, W, cam["tanfovx"], cam["tanfovy"], bg, 1.0, cam["viewmatrix"], cam["projmatrix"], 3,
                                       cam["campos"], False, False)
    rast = GaussianRasterizer(rs)
    # P == 0: zero images (background not applied), empty radii, backward works (rasterize_points.cu:83)
    e3 = torch.zeros(0, 3, device=dev(), requires_grad=True)
    out = rast(means3D=e3, means2D=torch.zeros(0, 3, device=dev()), opacities=torch.zeros(0, 1, device=dev()),
               shs=torch.zeros(0, 16, 3, device=dev()), scales=torch.zeros(0, 3, device=dev()),
               rotations=torch.zeros(0, 4, device=dev()))
    assert len(out) == 4 and out[0].shape == (3, H, W) and out[1].shape == (0,) and not out[0].any()
    # every Gaussian behind the near plane: R == 0, background everywhere, zero gradients
    m = scene["means3D"].clone()
    m[:, 2] = -1.0
    m.requires_grad_(True)
    sh = scene["shs"].clone().requires_grad_(True)
    color, radii, depth, alpha = rast(means3D=m, means2D=torch.zeros_like(m, requires_grad=True), opacities=scene["opacities"],
                                      shs=sh, scales=scene["scales"], rotations=scene["rotations"])
    assert not radii.any() and not alpha.any() and not depth.any()
    assert torch.allclose(color, bg[:, None, None].expand_as(color))
    (color.sum() + depth.sum() + alpha.sum()).backward()
    assert not m.grad.any() and not sh.grad.any()
    # markVisible == view-space z > 0.2, checked against the oracle
    vis = rast.markVisible(scene["means3D"])
    assert vis.dtype == torch.bool
    assert np.array_equal(np_(vis), golden.mark_visible(scene["means3D"], cam["viewmatrix"], cam["projmatrix"]))
    # non-contiguous / strided inputs are made contiguous like rasterize_points.cu:96-115
    big = torch.zeros(cfg["P"], 6, device=dev())
    big[:, ::2] = scene["means3D"]
    a = rast(means3D=big[:, ::2], means2D=torch.zeros_like(scene["means3D"]), opacities=scene["opacities"], shs=scene["shs"],
             scales=scene["scales"], rotations=scene["rotations"])
    b = rast(means3D=scene["means3D"], means2D=torch.zeros_like(scene["means3D"]), opacities=scene["opacities"],
             shs=scene["shs"], scales=scene["scales"], rotations=scene["rotations"])
    assert torch.equal(a[0], b[0]) and torch.equal(a[1], b[1])
    # debug=True path synchronises and returns the same result
    rsd = rs._replace(debug=True)
    c = GaussianRasterizer(rsd)(means3D=scene["means3D"], means2D=torch.zeros_like(scene["means3D"]), opacities=scene["opacities"],
                                shs=scene["shs"], scales=scene["scales"], rotations=scene["rotations"])
    assert torch.equal(c[0], b[0])
