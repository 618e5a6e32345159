"""Generates tests/golden/pyref_vectors.npz by IMPORTING the reference's own Python utilities from
/root/reference (cannot travel to the GPU box, hence the committed vectors):
  utils/sh_utils.eval_sh                      -> colours of the Python SH path
                                                 (scene/gaussian_model_ht.py:846-862)
  utils/general_utils.build_scaling_rotation,
  strip_symmetric                             -> 3D covariance of get_covariance
                                                 (scene/gaussian_model_ht.py:49-56)
  utils/graphics_utils.getWorld2View2,
  getProjectionMatrix                         -> camera matrices of scene/cameras.py:94-98
Run:  python tests/golden/make_golden_pyref.py      (in the build container)
The reference helpers hard-code device="cuda"; torch.zeros is shimmed to drop that argument while
they run -- the reference sources themselves are untouched."""
import importlib
import math
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
importlib.import_module("3dgs_hierarchical_training_b200").install()
import htgs_synthetic as syn  # noqa: E402

sys.path.insert(0, "/root/reference")
from utils import general_utils, graphics_utils, sh_utils  # noqa: E402

_zeros = torch.zeros


def _cpu_zeros(*a, **k):
    k.pop("device", None)
    return _zeros(*a, **k)


def main():
    P, W, H = 96, 64, 48
    scene = syn.make_scene(P, W, H, s_max=5.0, seed=77)
    cam = syn.make_camera(W, H, yaw_deg=5.0, pivot_z=3.0)
    out = {k: v.numpy() for k, v in scene.items()}
    # --- SH colours, exactly the expression of CF3DGS_Render.render's python path ---
    campos = cam["campos"]
    shs_view = scene["shs"].transpose(1, 2).view(-1, 3, 16)
    dir_pp = scene["means3D"] - campos.repeat(P, 1)
    dir_pp_normalized = dir_pp / dir_pp.norm(dim=1, keepdim=True)
    for deg in range(4):
        sh2rgb = sh_utils.eval_sh(deg, shs_view[:, :, :(deg + 1) ** 2], dir_pp_normalized)
        out[f"colors_deg{deg}"] = torch.clamp_min(sh2rgb + 0.5, 0.0).numpy()
    # --- covariance ---
    torch.zeros = _cpu_zeros
    general_utils.torch.zeros = _cpu_zeros
    try:
        for mod in (1.0, 0.6):
            L = general_utils.build_scaling_rotation(mod * scene["scales"], scene["rotations"])
            out[f"cov3D_mod{mod}"] = general_utils.strip_symmetric(L @ L.transpose(1, 2)).numpy()
    finally:
        torch.zeros = _zeros
        general_utils.torch.zeros = _zeros
    # --- camera matrices the way scene/cameras.py builds them (non-co3d branch) ---
    w2c = cam["viewmatrix"].t().double().numpy()  # our stored-transposed view -> standard [R|t]
    Rm, T = w2c[:3, :3].T, w2c[:3, 3]  # Camera stores R transposed (getWorld2View2 transposes it back)
    fovx = syn.FOVX_FRANCIS
    fovy = 2 * math.atan(cam["tanfovy"])
    wv = torch.tensor(graphics_utils.getWorld2View2(Rm, T)).transpose(0, 1)
    pm = graphics_utils.getProjectionMatrix(znear=0.01, zfar=100.0, fovX=fovx, fovY=fovy).transpose(0, 1)
    full = (wv.unsqueeze(0).bmm(pm.unsqueeze(0))).squeeze(0)
    out["ref_world_view_transform"] = wv.numpy()
    out["ref_full_proj_transform"] = full.numpy()
    out["ref_camera_center"] = wv.inverse()[3, :3].numpy()
    out["cam_meta"] = np.array([W, H, 5.0, 3.0], np.float64)
    np.savez_compressed(os.path.join(HERE, "pyref_vectors.npz"), **out)
    print("wrote", os.path.join(HERE, "pyref_vectors.npz"), {k: v.shape for k, v in out.items()})


if __name__ == "__main__":
    main()
