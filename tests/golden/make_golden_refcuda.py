"""Generates tests/golden/refcuda_*.npz on a GPU box by running the UNMODIFIED reference CUDA
rasterizer / simple-knn (oracle/_ref, built from /root/reference by oracle/Makefile) on seeded
synthetic scenes.  These are the golden vectors that pin the CPU oracle (and, transitively, every
parity test) to the reference itself: the reference ships none of its own.

Run on the GPU box:   python tests/golden/make_golden_refcuda.py gpurun_out/golden
then copy gpurun_out/golden/*.npz into tests/golden/ and commit."""
import importlib
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
importlib.import_module("3dgs_hierarchical_training_b200").install()
import htgs_synthetic as syn  # noqa: E402
from helpers import cov3d_of  # noqa: E402
from oracle import refcuda  # noqa: E402

CASES = {
    # name: (config, sh_degree, background, colors_precomp?, cov3D_precomp?, scale_modifier)
    "tiny_sh3": ("tiny", 3, (0.3, 0.6, 0.9), False, False, 1.0),
    "tiny_sh1_black": ("tiny", 1, (0.0, 0.0, 0.0), False, False, 1.0),
    "tiny_precomp_cov": ("tiny", 3, (0.1, 0.2, 0.3), True, True, 1.0),
    "ragged_sh3_mod": ("ragged", 3, (0.0, 0.0, 0.0), False, False, 0.8),
}


def main(outdir):
    os.makedirs(outdir, exist_ok=True)
    dev = torch.device("cuda:0")
    for name, (cfgname, deg, bgv, precomp, use_cov, mod) in CASES.items():
        cfg, scene, cam, up = syn.make_config(cfgname, device=dev)
        P = cfg["P"]
        bg = torch.tensor(bgv, dtype=torch.float32, device=dev)
        colors = torch.rand(P, 3, generator=torch.Generator().manual_seed(5)).to(dev) if precomp else None
        cov = cov3d_of({k: v.cpu() for k, v in scene.items()}).to(dev) if use_cov else None
        kw = dict(shs=None if precomp else scene["shs"], colors_precomp=colors, scales=None if use_cov else scene["scales"],
                  rotations=None if use_cov else scene["rotations"], cov3D_precomp=cov, scale_modifier=mod)
        fw = refcuda.forward(scene["means3D"], scene["opacities"], cam, bg, deg, **kw)
        g = refcuda.backward(fw, scene["means3D"], cam, bg, up["dL_dcolor"], up["dL_ddepth"], up["dL_dalpha"], deg, **kw)
        torch.cuda.synchronize()
        out = {"case": np.array([cfgname, str(deg), str(precomp), str(use_cov), str(mod)]), "bg": np.array(bgv, np.float32),
               "R": np.array([fw["R"]], np.int64)}
        for k in ("color", "depth", "alpha", "radii", "depths", "clamped", "means2D", "cov3D", "conic_opacity", "rgb",
                  "tiles_touched", "point_offsets", "point_list", "point_list_keys", "ranges", "n_contrib"):
            out[k] = fw[k].cpu().numpy()
        for k, v in g.items():
            out[k] = v.cpu().numpy()
        if colors is not None:
            out["colors_precomp"] = colors.cpu().numpy()
        if cov is not None:
            out["cov3D_precomp"] = cov.cpu().numpy()
        np.savez_compressed(os.path.join(outdir, f"refcuda_{name}.npz"), **out)
        print(name, "R", fw["R"], "visible", int((fw["radii"] > 0).sum()))
    # simple-knn
    pts = syn.make_scene(3000, 64, 64, seed=91)["means3D"].to(dev)
    d = refcuda.dist2(pts)
    vis = refcuda.mark_visible(pts, syn.make_camera(64, 64, device=dev)["viewmatrix"], syn.make_camera(64, 64, device=dev)["projmatrix"])
    np.savez_compressed(os.path.join(outdir, "refcuda_knn.npz"), points=pts.cpu().numpy(), dist2=d.cpu().numpy(),
                        visible=vis.cpu().numpy())
    print("knn ok")


if __name__ == "__main__":
    main(sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "gpurun_out", "golden"))
