"""Stage-by-stage diagnostics on a GPU box: ours vs the CPU oracle vs the reference CUDA build.
Development aid (prints everything, never asserts); the real gates are tests/ -m gpu.
Usage: python tools/gpu_check.py [config ...]"""
import importlib
import os
import sys
import time
import traceback

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
pkg = importlib.import_module("3dgs_hierarchical_training_b200")
pkg.install()
import htgs_native as nat  # noqa: E402
import htgs_synthetic as syn  # noqa: E402
from diff_gaussian_rasterization import GaussianRasterizationSettings, GaussianRasterizer  # noqa: E402
from diff_gaussian_rasterization import _C  # noqa: E402
from oracle import golden, refcuda  # noqa: E402

dev = torch.device("cuda:0")


def view_state(geom, binning, img, P, R, W, H):
    cap = nat.binning_capacity_from_bytes(binning.numel(), W, H) if R > 0 else 0
    gl, bl, il = nat.geom_layout(P), nat.binning_layout(cap, W, H), nat.image_layout(W, H)
    tiles = ((W + 15) // 16) * ((H + 15) // 16)

    def sl(buf, off, n, dtype):
        nbytes = n * torch.empty((), dtype=dtype).element_size()
        return buf[off:off + nbytes].view(dtype)
    st = dict(means2D=sl(geom, gl["means2D"], 2 * P, torch.float32).view(P, 2),
              conic_opacity=sl(geom, gl["conic_opacity"], 4 * P, torch.float32).view(P, 4),
              feat=sl(geom, gl["feat"], 4 * P, torch.float32).view(P, 4),
              clamped=sl(geom, gl["clamped"], P, torch.uint8),
              tiles_touched=sl(geom, gl["tiles_touched"], P, torch.int32),
              point_offsets=sl(geom, gl["point_offsets"], P, torch.int32),
              ranges=sl(img, il["ranges"], 2 * tiles, torch.int32).view(tiles, 2),
              n_contrib=sl(img, il["n_contrib"], W * H, torch.int32).view(H, W))
    if R > 0:
        st["keys"] = sl(binning, bl["keys_sorted"], R, torch.int64)
        st["point_list"] = sl(binning, bl["values_sorted"], R, torch.int32)
    return st


def cmp(name, a, b, exact=False):
    a = a.detach().cpu().numpy() if hasattr(a, "detach") else np.asarray(a)
    b = b.detach().cpu().numpy() if hasattr(b, "detach") else np.asarray(b)
    if a.shape != b.shape:
        print(f"  {name}: SHAPE {a.shape} vs {b.shape}")
        return
    if a.size == 0:
        print(f"  {name}: empty")
        return
    if exact:
        nbad = int((a.view(np.uint8) != b.view(np.uint8)).reshape(a.shape + (-1,)).any(-1).sum()) if a.dtype.kind == "f" else int((a != b).sum())
        print(f"  {name}: mismatches {nbad}/{a.size}")
    else:
        a64, b64 = a.astype(np.float64), b.astype(np.float64)
        d = np.abs(a64 - b64)
        ref = np.abs(b64).max()
        print(f"  {name}: max|d| {d.max():.3e}  rel(inf) {d.max() / max(ref, 1e-30):.3e}  ref max {ref:.3e} n>1e-5 {(d > 1e-5).sum()}")


def run(name, bgv=(0.3, 0.6, 0.9), colors_precomp=False, deg=3):
    cfg, scene, cam, up = syn.make_config(name, device=dev)
    P, W, H = cfg["P"], cfg["W"], cfg["H"]
    bg = torch.tensor(bgv, dtype=torch.float32, device=dev)
    print(f"== {name} P={P} {W}x{H} deg={deg} precomp={colors_precomp}")
    kw = dict(scales=scene["scales"], rotations=scene["rotations"])
    colors = torch.rand(P, 3, generator=torch.Generator().manual_seed(5)).to(dev) if colors_precomp else None
    shs = None if colors_precomp else scene["shs"]
    empty = torch.Tensor([])
    args = (bg, scene["means3D"], colors if colors_precomp else empty, scene["opacities"], scene["scales"], scene["rotations"],
            1.0, empty, cam["viewmatrix"], cam["projmatrix"], cam["tanfovx"], cam["tanfovy"], H, W,
            shs if shs is not None else empty, deg, cam["campos"], False, True)
    t0 = time.time()
    R, color, depth, alpha, radii, geom, binning, img = _C.rasterize_gaussians(*args)
    torch.cuda.synchronize()
    print(f"  ours fwd ok R={R} t={time.time()-t0:.3f}s")
    mine = view_state(geom, binning, img, P, R, W, H)
    gb = _C.rasterize_gaussians_backward(bg, scene["means3D"], radii, colors if colors_precomp else empty, scene["scales"],
                                         scene["rotations"], 1.0, empty, cam["viewmatrix"], cam["projmatrix"], cam["tanfovx"],
                                         cam["tanfovy"], up["dL_dcolor"], up["dL_ddepth"], up["dL_dalpha"],
                                         shs if shs is not None else empty, deg, cam["campos"], geom, R, binning, img, alpha, True)
    torch.cuda.synchronize()
    names = ["dL_dmeans2D", "dL_dcolors", "dL_dopacity", "dL_dmeans3D", "dL_dcov3D", "dL_dsh", "dL_dscales", "dL_drotations"]
    gmine = dict(zip(names, gb))
    print("  ours bwd ok")

    if P <= 400_000:
        t0 = time.time()
        o = golden.forward(scene["means3D"], scene["opacities"], cam, bg, deg, shs=shs, colors_precomp=colors, **kw)
        og = golden.backward(o, up["dL_dcolor"], up["dL_ddepth"], up["dL_dalpha"])
        print(f"  -- vs CPU oracle (R={o['R']}, {time.time()-t0:.2f}s, tested={o['counters'][0]} blended={o['counters'][1]})")
        vis = o["radii"] > 0
        cmp("radii", radii, o["radii"], exact=True)
        cmp("tiles_touched", mine["tiles_touched"], o["tiles_touched"].astype(np.int32), exact=True)
        cmp("point_offsets", mine["point_offsets"], o["point_offsets"].astype(np.int32), exact=True)
        cmp("means2D[vis]", mine["means2D"].cpu().numpy()[vis], o["means2D"][vis], exact=True)
        cmp("conic_opacity[vis]", mine["conic_opacity"].cpu().numpy()[vis], o["conic_opacity"][vis], exact=True)
        cmp("depth[vis]", mine["feat"].cpu().numpy()[vis, 3], o["depths"][vis], exact=True)
        if not colors_precomp:
            cmp("rgb[vis]", mine["feat"].cpu().numpy()[vis, :3], o["rgb"][vis], exact=True)
        if R > 0 and R == o["R"]:
            cmp("keys", mine["keys"], o["keys"].astype(np.int64), exact=True)
            cmp("point_list", mine["point_list"], o["point_list"].astype(np.int32), exact=True)
        cmp("ranges", mine["ranges"], o["ranges"].astype(np.int32), exact=True)
        cmp("n_contrib", mine["n_contrib"], o["n_contrib"].astype(np.int32), exact=True)
        cmp("color", color, o["color"]); cmp("depth", depth, o["depth"]); cmp("alpha", alpha, o["alpha"])
        for k in names:
            if k == "dL_dcolors" and not colors_precomp:
                continue
            if k == "dL_dsh" and colors_precomp:
                continue
            cmp(k, gmine[k], og[k])

    if refcuda.available():
        r = refcuda.forward(scene["means3D"], scene["opacities"], cam, bg, deg, shs=shs, colors_precomp=colors, **kw)
        rg = refcuda.backward(r, scene["means3D"], cam, bg, up["dL_dcolor"], up["dL_ddepth"], up["dL_dalpha"], deg, shs=shs,
                              colors_precomp=colors, **kw)
        torch.cuda.synchronize()
        print(f"  -- vs reference CUDA (R={r['R']})")
        vis = (r["radii"] > 0)
        cmp("radii", radii, r["radii"], exact=True)
        cmp("tiles_touched", mine["tiles_touched"], r["tiles_touched"], exact=True)
        cmp("point_offsets", mine["point_offsets"], r["point_offsets"], exact=True)
        cmp("means2D[vis]", mine["means2D"][vis], r["means2D"].view(P, 2)[vis], exact=True)
        cmp("conic_opacity[vis]", mine["conic_opacity"][vis], r["conic_opacity"].view(P, 4)[vis], exact=True)
        cmp("depth[vis]", mine["feat"][vis, 3], r["depths"][vis], exact=True)
        if not colors_precomp:
            cmp("rgb[vis]", mine["feat"][vis, :3], r["rgb"].view(P, 3)[vis], exact=True)
            cmp("clamped[vis]", ((mine["clamped"][:, None] >> torch.arange(3, device=dev)[None]) & 1)[vis].to(torch.uint8),
                r["clamped"].view(P, 3)[vis], exact=True)
        if R > 0 and R == r["R"]:
            cmp("keys", mine["keys"], r["point_list_keys"], exact=True)
            cmp("point_list", mine["point_list"], r["point_list"], exact=True)
        tiles = mine["ranges"].shape[0]
        cmp("ranges", mine["ranges"], r["ranges"].view(-1, 2)[:tiles], exact=True)
        cmp("n_contrib", mine["n_contrib"], r["n_contrib"].view(H, W), exact=True)
        cmp("color(bits)", color, r["color"], exact=True)
        cmp("depth(bits)", depth, r["depth"], exact=True)
        cmp("alpha(bits)", alpha, r["alpha"], exact=True)
        cmp("color", color, r["color"]); cmp("depth", depth, r["depth"]); cmp("alpha", alpha, r["alpha"])
        # reference self-noise: run its backward twice
        r2 = refcuda.forward(scene["means3D"], scene["opacities"], cam, bg, deg, shs=shs, colors_precomp=colors, **kw)
        rg2 = refcuda.backward(r2, scene["means3D"], cam, bg, up["dL_dcolor"], up["dL_ddepth"], up["dL_dalpha"], deg, shs=shs,
                               colors_precomp=colors, **kw)
        for k in names:
            if k == "dL_dcolors" and not colors_precomp:
                continue
            if k == "dL_dsh" and colors_precomp:
                continue
            cmp(k, gmine[k], rg[k].view(gmine[k].shape))
            cmp(k + " (ref self-noise)", rg2[k], rg[k])
    return True


if __name__ == "__main__":
    names = sys.argv[1:] or ["tiny", "small", "ragged", "deep", "config1"]
    for n in names:
        try:
            if n.endswith(":precomp"):
                run(n.split(":")[0], colors_precomp=True)
            elif ":deg" in n:
                run(n.split(":")[0], deg=int(n.split(":deg")[1]))
            else:
                run(n)
        except Exception:
            traceback.print_exc()
            print(f"!! {n} FAILED")
            try:
                torch.cuda.synchronize()
            except Exception:
                print("!! CUDA context is dead; stopping")
                break
