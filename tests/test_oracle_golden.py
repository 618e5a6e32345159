"""CPU tests: the oracle against the committed golden vectors.

  pyref_vectors.npz   produced by importing the reference's Python utilities
                      (tests/golden/make_golden_pyref.py): SH colours, 3D covariance, camera matrices.
  refcuda_*.npz       produced by running the UNMODIFIED reference CUDA code on a B200
                      (tests/golden/make_golden_refcuda.py): every stage output and all gradients.

Bars (BASELINE.json): integer/index outputs bit-exact; images <= 1e-5 max-abs; gradients <= 1e-4
relative.  The only allowance is documented in helpers.ORACLE_PIXEL_OUTLIER_FRAC: the CPU cannot
reproduce the GPU's ex2.approx bit for bit, so a pair sitting exactly on one of the reference's hard
thresholds may flip in the oracle."""
import glob
import math
import os

import numpy as np
import pytest
import torch

import htgs_synthetic as syn
from helpers import GRAD_RTOL, IMAGE_ATOL, ORACLE_GRAD_RTOL, ORACLE_PIXEL_OUTLIER_FRAC, cov3d_of, rel_inf
from oracle import golden

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def _pyref():
    return np.load(os.path.join(GOLD, "pyref_vectors.npz"))


@pytest.mark.parametrize("deg", [0, 1, 2, 3])
def test_sh_colours_match_reference_eval_sh(deg):
    v = _pyref()
    W, H, yaw, pivot = v["cam_meta"]
    cam = syn.make_camera(int(W), int(H), yaw_deg=float(yaw), pivot_z=float(pivot))
    st = golden.forward(v["means3D"], v["opacities"], cam, np.zeros(3, np.float32), deg, shs=v["shs"], scales=v["scales"],
                        rotations=v["rotations"])
    vis = st["radii"] > 0
    assert vis.sum() > 40
    assert np.abs(st["rgb"][vis] - v[f"colors_deg{deg}"][vis]).max() <= 2e-6
    assert np.array_equal(st["clamped"][vis].astype(bool), (v[f"colors_deg{deg}"][vis] == 0))


@pytest.mark.parametrize("mod", [1.0, 0.6])
def test_cov3d_matches_reference_build_scaling_rotation(mod):
    v = _pyref()
    W, H, yaw, pivot = v["cam_meta"]
    cam = syn.make_camera(int(W), int(H), yaw_deg=float(yaw), pivot_z=float(pivot))
    st = golden.forward(v["means3D"], v["opacities"], cam, np.zeros(3, np.float32), 0, shs=v["shs"], scales=v["scales"],
                        rotations=v["rotations"], scale_modifier=mod)
    vis = st["radii"] > 0
    ref = v[f"cov3D_mod{mod}"]
    assert np.abs(st["cov3D"][vis] - ref[vis]).max() <= 1e-6 * np.abs(ref[vis]).max()
    # and the test-side helper used for the cov3D_precomp path agrees with the reference too
    mine = cov3d_of({k: torch.from_numpy(v[k]) for k in ("rotations", "scales")}).numpy()
    assert np.abs(mine - v["cov3D_mod1.0"]).max() <= 1e-6 * np.abs(v["cov3D_mod1.0"]).max()


def test_synthetic_camera_follows_reference_conventions():
    v = _pyref()
    W, H, yaw, pivot = v["cam_meta"]
    cam = syn.make_camera(int(W), int(H), yaw_deg=float(yaw), pivot_z=float(pivot))
    assert np.abs(cam["viewmatrix"].numpy() - v["ref_world_view_transform"]).max() <= 1e-6
    assert np.abs(cam["projmatrix"].numpy() - v["ref_full_proj_transform"]).max() <= 2e-5
    assert np.abs(cam["campos"].numpy() - v["ref_camera_center"]).max() <= 1e-5


REFCUDA = sorted(glob.glob(os.path.join(GOLD, "refcuda_*.npz")))
REFCUDA_CASES = [p for p in REFCUDA if not p.endswith("refcuda_knn.npz")]


def test_refcuda_goldens_present():
    assert len(REFCUDA_CASES) >= 3, "run tests/golden/make_golden_refcuda.py on a GPU box and commit the vectors"


@pytest.mark.parametrize("path", REFCUDA_CASES, ids=[os.path.basename(p)[8:-4] for p in REFCUDA_CASES])
def test_oracle_matches_reference_cuda(path):
    v = np.load(path)
    cfgname, deg, precomp, use_cov, mod = v["case"]
    deg, precomp, use_cov, mod = int(deg), precomp == "True", use_cov == "True", float(mod)
    cfg, scene, cam, up = syn.make_config(str(cfgname))
    P, W, H = cfg["P"], cfg["W"], cfg["H"]
    kw = dict(shs=None if precomp else scene["shs"], colors_precomp=v["colors_precomp"] if precomp else None,
              scales=None if use_cov else scene["scales"], rotations=None if use_cov else scene["rotations"],
              cov3D_precomp=v["cov3D_precomp"] if use_cov else None, scale_modifier=mod)
    st = golden.forward(scene["means3D"], scene["opacities"], cam, v["bg"], deg, **kw)
    vis = v["radii"] > 0
    # ---- integer / index outputs: bit-exact ----
    assert np.array_equal(st["radii"], v["radii"])
    assert np.array_equal(st["tiles_touched"], v["tiles_touched"].view(np.uint32))
    assert np.array_equal(st["point_offsets"], v["point_offsets"].view(np.uint32))
    assert st["R"] == int(v["R"][0])
    assert np.array_equal(st["keys"], v["point_list_keys"].view(np.uint64))
    assert np.array_equal(st["point_list"], v["point_list"].view(np.uint32))
    tiles = st["ranges"].shape[0]
    assert np.array_equal(st["ranges"], v["ranges"].view(np.uint32).reshape(-1, 2)[:tiles])
    # ---- per-Gaussian float state: bit-exact too (same fp32 operation order as the SASS) ----
    assert np.array_equal(st["depths"][vis].view(np.uint32), v["depths"][vis].view(np.uint32))
    assert np.array_equal(st["means2D"][vis].view(np.uint32), v["means2D"].reshape(P, 2)[vis].view(np.uint32))
    assert np.array_equal(st["conic_opacity"][vis].view(np.uint32), v["conic_opacity"].reshape(P, 4)[vis].view(np.uint32))
    if not precomp:
        assert np.array_equal(st["rgb"][vis].view(np.uint32), v["rgb"].reshape(P, 3)[vis].view(np.uint32))
        assert np.array_equal(st["clamped"][vis], v["clamped"].reshape(P, 3)[vis])
    if not use_cov:
        assert np.array_equal(st["cov3D"][vis].view(np.uint32), v["cov3D"].reshape(P, 6)[vis].view(np.uint32))
    # ---- images ----
    n_pix = W * H
    budget = max(1, int(math.ceil(ORACLE_PIXEL_OUTLIER_FRAC * n_pix)))
    assert (st["n_contrib"] != v["n_contrib"].view(np.uint32).reshape(H, W)).sum() <= budget
    for k in ("color", "depth", "alpha"):
        bad = np.abs(st[k] - v[k]) > IMAGE_ATOL * max(1.0, np.abs(v[k]).max())
        assert bad.sum() <= budget * st[k].shape[0], f"{k}: {bad.sum()} values beyond {IMAGE_ATOL}"
    # ---- gradients ----
    g = golden.backward(st, up["dL_dcolor"], up["dL_ddepth"], up["dL_dalpha"])
    flipped = (st["n_contrib"] != v["n_contrib"].view(np.uint32).reshape(H, W)).any() or any(
        (np.abs(st[k] - v[k]) > IMAGE_ATOL * max(1.0, np.abs(v[k]).max())).any() for k in ("color", "depth", "alpha"))
    tol = ORACLE_GRAD_RTOL if flipped else GRAD_RTOL
    names = ["dL_dmeans2D", "dL_dopacity", "dL_dmeans3D", "dL_dcov3D", "dL_dconic", "dL_ddepths"]
    names += ["dL_dcolors"] if precomp else ["dL_dsh"]
    names += [] if use_cov else ["dL_dscales", "dL_drotations"]
    for k in names:
        ref = v[k].reshape(g[k].shape)
        assert rel_inf(g[k], ref) <= tol, f"{k}: rel err {rel_inf(g[k], ref):.3e} > {tol}"


def test_oracle_knn_and_visibility_match_reference_cuda():
    p = os.path.join(GOLD, "refcuda_knn.npz")
    if not os.path.exists(p):
        pytest.skip("refcuda_knn.npz not generated yet")
    v = np.load(p)
    d = golden.dist2(v["points"])
    assert np.array_equal(d.view(np.uint32), v["dist2"].view(np.uint32)), "distCUDA2 must be bit-exact"
    cam = syn.make_camera(64, 64)
    assert np.array_equal(golden.mark_visible(v["points"], cam["viewmatrix"], cam["projmatrix"]), v["visible"])


def test_oracle_sort_and_ranges_properties():
    """Size-independent properties at a larger size: keys sorted, stable on ties, ranges partition the list."""
    cfg, scene, cam, up = syn.make_config("config1")
    st = golden.forward(scene["means3D"], scene["opacities"], cam, np.zeros(3, np.float32), 3, shs=scene["shs"],
                        scales=scene["scales"], rotations=scene["rotations"])
    k, pl = st["keys"], st["point_list"]
    assert (k[1:] >= k[:-1]).all()
    tie = k[1:] == k[:-1]
    assert (pl[1:][tie] > pl[:-1][tie]).all()  # equal (tile, depth): Gaussian index order = stable sort
    assert np.array_equal(np.sort(st["values_unsorted"]), np.sort(pl))
    assert st["point_offsets"][-1] == st["R"] == st["tiles_touched"].sum()
    r = st["ranges"]
    nonempty = r[:, 1] > r[:, 0]
    assert (r[nonempty, 1] - r[nonempty, 0]).sum() == st["R"]
    tile_of = (k >> np.uint64(32)).astype(np.int64)
    for t in np.flatnonzero(nonempty)[::17]:
        assert (tile_of[r[t, 0]:r[t, 1]] == t).all()
    assert golden.higher_msb(256) == 9 and golden.higher_msb(8160) == 13 and golden.higher_msb(2040) == 11


def test_oracle_edge_cases():
    cfg, scene, cam, up = syn.make_config("tiny")
    bg = np.array([0.2, 0.4, 0.6], np.float32)
    # P == 0: images stay zero, background NOT applied (rasterize_points.cu:83)
    st = golden.forward(np.zeros((0, 3), np.float32), np.zeros((0, 1), np.float32), cam, bg, 3)
    assert st["R"] == 0 and not st["color"].any()
    # everything behind the near plane: every pixel is the background, n_contrib 0
    m = scene["means3D"].clone()
    m[:, 2] = -1.0
    st = golden.forward(m, scene["opacities"], cam, bg, 3, shs=scene["shs"], scales=scene["scales"], rotations=scene["rotations"])
    assert st["R"] == 0 and (st["radii"] == 0).all() and not st["n_contrib"].any()
    assert np.allclose(st["color"], bg[:, None, None]) and not st["alpha"].any() and not st["depth"].any()
    g = golden.backward(st, up["dL_dcolor"], up["dL_ddepth"], up["dL_dalpha"])
    assert all(not v.any() for v in g.values())
