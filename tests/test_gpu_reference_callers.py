"""The reference's OWN callers of the plugin, unchanged, running on this repository's rasterizer.

north_star: "... so gaussian_renderer/__init__.py render() and the HT-3DGS trainer (trainer/ht3dgs_trainer.py,
scene/gaussian_model_ht.py) use it unchanged".  tools/stage_reference_callers.py copies the reference's
gaussian_renderer/, scene/, utils/ and arguments/ packages byte for byte into baseline/_ref/htgs_callers/ (git-ignored,
travels to the GPU box); tests/ref_callers.py imports them with this repository's `diff_gaussian_rasterization` and
`simple_knn._C` on sys.path and stubs only the third-party packages the image lacks (lietorch, plyfile, matplotlib, ...).

Covered here, through the reference's code and not a re-typed imitation of it:
  * gaussian_renderer.render()                    gaussian_renderer/__init__.py:19-113 (SH path, convert_SHs_python path,
                                                  compute_cov3D_python path, override_color)
  * CF3DGS_Render.init_model -> HTGaussianModel.create_from_pcd -> distCUDA2       scene/gaussian_model_ht.py:197-233, :742-770
  * CF3DGS_Render.render()                        scene/gaussian_model_ht.py:775-908 (the `len(out) == 4` branch, clamp, dict)
  * scene.cameras.Camera                          scene/cameras.py:16-98 (the matrices the kernels consume)
  * a train_step-shaped loop: loss.backward(), viewspace_points.grad, add_densification_stats, optimizer.step
                                                  trainer/ht3dgs_trainer.py:137-166, scene/gaussian_model_ht.py:275-289, :718-721
Results are checked against the CPU oracle."""
import math
import os
import sys
import types

import numpy as np
import pytest
import torch

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import ref_callers  # noqa: E402
from helpers import GRAD_RTOL, IMAGE_ATOL, ORACLE_GRAD_RTOL, np_, rel_inf  # noqa: E402
from oracle import golden  # noqa: E402

pytestmark = [pytest.mark.gpu,
              pytest.mark.skipif(not ref_callers.staged(), reason="baseline/_ref/htgs_callers not staged (tools/stage_reference_callers.py)")]

W, H = 160, 96
FOVX = 1.3541787529604106


@pytest.fixture(scope="module")
def ref():
    return ref_callers.load()


def _camera(ref, uid=0, yaw=0.0):
    """A scene.cameras.Camera of the reference (not a stand-in): R, T as the dataset readers produce them."""
    fovy = 2.0 * math.atan(math.tan(FOVX / 2.0) * H / W)
    c, s = math.cos(yaw), math.sin(yaw)
    R = np.array([[c, 0.0, s], [0.0, 1.0, 0.0], [-s, 0.0, c]], dtype=np.float64)
    T = np.array([0.05, -0.02, 0.3], dtype=np.float64)
    K = np.array([[W / (2 * math.tan(FOVX / 2)), 0, W / 2], [0, H / (2 * math.tan(fovy / 2)), H / 2], [0, 0, 1]], dtype=np.float64)
    image = torch.rand(3, H, W, generator=torch.Generator().manual_seed(uid))
    return ref.cameras.Camera(colmap_id=uid, R=R, T=T, FoVx=FOVX, FoVy=fovy, image=image, gt_alpha_mask=None,
                              image_name=f"{uid:05d}", uid=uid, intrinsics=K, data_device="cuda")


def _point_cloud(ref, n=3000, seed=0):
    g = np.random.default_rng(seed)
    z = g.uniform(1.5, 6.0, n)
    x = z * math.tan(FOVX / 2) * g.uniform(-1.0, 1.0, n)
    y = z * math.tan(FOVX / 2) * H / W * g.uniform(-1.0, 1.0, n)
    return ref.graphics_utils.BasicPointCloud(points=np.stack([x, y, z], 1), colors=g.uniform(0, 1, (n, 3)), normals=np.zeros((n, 3)))


def _oracle(cam, means3D, opacities, bg, deg, **kw):
    camd = dict(W=W, H=H, tanfovx=math.tan(cam.FoVx * 0.5), tanfovy=math.tan(cam.FoVy * 0.5), viewmatrix=cam.world_view_transform,
                projmatrix=cam.full_proj_transform, campos=cam.camera_center)
    return golden.forward(means3D.detach(), opacities.detach(), camd, bg, deg, **{k: (v.detach() if v is not None else None) for k, v in kw.items()})


def _images_close(mine, want, what):
    bad = np.abs(np_(mine) - want) > IMAGE_ATOL * max(1.0, np.abs(want).max())
    assert bad.mean() <= 2e-4, f"{what}: {bad.sum()} values beyond tolerance"


def test_modules_resolve_to_this_repository(ref):
    import diff_gaussian_rasterization
    import simple_knn._C as knn
    here = os.path.join(ref_callers.ROOT, "3dgs_hierarchical_training_b200")
    assert os.path.dirname(os.path.dirname(diff_gaussian_rasterization.__file__)) == here
    assert ref.gaussian_renderer.GaussianRasterizer is diff_gaussian_rasterization.GaussianRasterizer
    assert ref.gaussian_model_ht.GaussianRasterizer is diff_gaussian_rasterization.GaussianRasterizer
    assert ref.gaussian_model_ht.distCUDA2 is knn.distCUDA2
    for m in (ref.gaussian_renderer, ref.gaussian_model_ht, ref.cameras):
        assert m.__file__.startswith(ref_callers.STAGED)
    # nothing on the rasterizer path is a stub
    assert "diff_gaussian_rasterization" not in ref.stubbed and "simple_knn" not in ref.stubbed


def test_cf3dgs_render_and_train_step_shape(ref):
    """CF3DGS_Render (the HT trainer's renderer): init from a point cloud through distCUDA2, render, loss, backward,
    densification statistics, Adam -- every call below is a method of the reference's classes."""
    M = ref.gaussian_model_ht
    cam = _camera(ref)
    r = M.CF3DGS_Render(sh_degree=3, white_background=False, view_dependent=True)
    pcd = _point_cloud(ref)
    r.init_model(pcd)  # -> HTGaussianModel.create_from_pcd -> distCUDA2 (gaussian_model_ht.py:211-216)
    g = r.gaussians
    P = g._xyz.shape[0]
    assert P == 3000 and g._scaling.shape == (P, 3)
    # the initial scales really came from this library's distCUDA2: log(sqrt(clamp_min(dist2, 1e-7))), repeated x3
    d2 = golden.dist2(torch.from_numpy(np.asarray(pcd.points)).float())
    assert np.allclose(np_(g._scaling)[:, 0], np.log(np.sqrt(np.maximum(d2, 1e-7))), rtol=1e-5, atol=1e-6)
    g._scaling.data += 0.8  # nearest-neighbour sized blobs are sub-pixel here; make them cover a few pixels
    g._features_rest.data.normal_(0.0, 0.2, generator=None)
    g.active_sh_degree = 3
    opt = types.SimpleNamespace(percent_dense=0.01, position_lr_init=1.6e-4, position_lr_final=1.6e-6, position_lr_delay_mult=0.01,
                                position_lr_max_steps=30000, feature_lr=2.5e-3, opacity_lr=0.05, scaling_lr=5e-3, rotation_lr=1e-3)
    g.training_setup(opt)

    pkg = r.render(cam)  # gaussian_model_ht.py:775-908
    assert set(pkg) == {"image", "depth", "alpha", "viewspace_points", "visibility_filter", "radii"}  # the len(out) == 4 branch
    assert pkg["image"].shape == (3, H, W) and pkg["depth"].shape == (1, H, W) and pkg["alpha"].shape == (1, H, W)
    assert pkg["radii"].dtype == torch.int32 and pkg["visibility_filter"].dtype == torch.bool
    assert int(pkg["visibility_filter"].sum()) > P // 2
    o = _oracle(cam, g.get_xyz, g.get_opacity, np_(r.bg_color), 3, shs=g.get_features, scales=g.get_scaling, rotations=g.get_rotation)
    assert np.array_equal(np_(pkg["radii"]), o["radii"])
    _images_close(pkg["image"], np.clip(o["color"], 0.0, 1.0), "image")
    _images_close(pkg["depth"], o["depth"], "depth")
    _images_close(pkg["alpha"], o["alpha"], "alpha")

    # train_step-shaped tail (ht3dgs_trainer.py:127-166): in-place edit of the returned depth, loss, backward,
    # max_radii2D / densification statistics on the visible set, optimizer step
    depth = pkg["depth"]
    depth[depth < 0.01] = 0.01  # compute_loss edits the plugin's output in place (ht3dgs_trainer.py:1292)
    loss = (pkg["image"] - cam.original_image).abs().mean() + 0.01 * depth.mean() + 0.01 * pkg["alpha"].mean()
    loss.backward()
    vs, vis, radii = pkg["viewspace_points"], pkg["visibility_filter"], pkg["radii"]
    assert vs.grad is not None and vs.grad.shape == (P, 3) and bool(torch.isfinite(vs.grad).all())
    assert not vs.grad[~vis].any() and bool(vs.grad[vis, :2].abs().sum() > 0) and not vs.grad[:, 2].any()
    for p in (g._xyz, g._features_dc, g._features_rest, g._scaling, g._rotation, g._opacity):
        assert p.grad is not None and bool(torch.isfinite(p.grad).all()) and bool(p.grad.abs().sum() > 0)
    g.max_radii2D[vis] = torch.max(g.max_radii2D[vis], radii[vis])
    g.add_densification_stats(vs, vis)  # gaussian_model_ht.py:718-721
    assert torch.equal(g.denom[:, 0] > 0, vis)
    before = g._xyz.detach().clone()
    g.optimizer.step()
    g.optimizer.zero_grad(set_to_none=True)
    assert not torch.equal(before, g._xyz.detach())
    # a second render after the update still works and differs
    again = r.render(cam)
    assert not torch.equal(again["image"], pkg["image"])


def test_cf3dgs_render_python_paths(ref):
    """convert_SHs_python (colors_precomp built by the reference's eval_sh) and compute_cov3D_python (cov3D_precomp
    built by the reference's build_scaling_rotation) must give the image of the native SH / scale+rotation path."""
    M = ref.gaussian_model_ht
    cam = _camera(ref, uid=1, yaw=0.05)
    r = M.CF3DGS_Render(sh_degree=3, white_background=True, view_dependent=True)
    r.init_model(_point_cloud(ref, seed=1))
    g = r.gaussians
    g._scaling.data += 0.7
    g._rotation.data.normal_()
    g._features_rest.data.normal_(0.0, 0.2)
    g.active_sh_degree = 2
    native = r.render(cam)
    py_sh = r.render(cam, convert_SHs_python=True)       # uses get_RT(uid): identity without pose parameters
    py_cov = r.render(cam, compute_cov3D_python=True)
    inv = r.render(cam, invert_bg_color=True)
    assert torch.equal(py_cov["radii"], native["radii"]) or float((py_cov["radii"] != native["radii"]).float().mean()) < 1e-3
    assert float((py_cov["image"] - native["image"]).abs().max()) < 2e-3  # cov3D from torch matmul vs in-kernel: last-bit radii
    # the Python SH path takes the camera centre from get_RT (identity here) rather than from the camera: compare with the
    # oracle fed the same precomputed colours instead of with the native path
    shs_view = g.get_features.transpose(1, 2).view(-1, 3, 16)
    dirs = g._xyz - torch.eye(4, device="cuda").inverse()[:3, 3][None]
    colors = torch.clamp_min(ref.sh_utils.eval_sh(2, shs_view, dirs / dirs.norm(dim=1, keepdim=True)) + 0.5, 0.0)
    o = _oracle(cam, g.get_xyz, g.get_opacity, np_(r.bg_color), 2, colors_precomp=colors, scales=g.get_scaling, rotations=g.get_rotation)
    _images_close(py_sh["image"], np.clip(o["color"], 0.0, 1.0), "convert_SHs_python image")
    assert not torch.equal(inv["image"], native["image"]) and torch.equal(inv["alpha"], native["alpha"])


def test_gaussian_renderer_render(ref):
    """The Inria-style render() of gaussian_renderer/__init__.py with a minimal `pc` exposing the getters it reads
    (the legacy GaussianModel's own get_xyz starts with `assert False`, scene/gaussian_model.py:152-153)."""
    import htgs_synthetic as syn
    cam = _camera(ref, uid=2, yaw=-0.04)
    scene = syn.make_scene(5000, W, H, s_max=5.0, seed=21, device="cuda")
    cov = ref.gaussian_model_ht.HTGaussianModel(3).covariance_activation
    leaves = {k: v.clone().requires_grad_(True) for k, v in scene.items()}
    pc = types.SimpleNamespace(get_xyz=leaves["means3D"], get_opacity=leaves["opacities"], get_scaling=leaves["scales"],
                               get_rotation=leaves["rotations"], get_features=leaves["shs"], active_sh_degree=3, max_sh_degree=3,
                               get_covariance=lambda m=1.0: cov(leaves["scales"], m, leaves["rotations"]))
    bg = torch.tensor([0.2, 0.1, 0.4], device="cuda")
    pipe = types.SimpleNamespace(debug=False, compute_cov3D_python=False, convert_SHs_python=False)
    out = ref.gaussian_renderer.render(cam, pc, pipe, bg)
    assert set(out) == {"render", "viewspace_points", "visibility_filter", "radii", "depth", "alpha"}  # len(render_out) > 2 branch
    o = _oracle(cam, leaves["means3D"], leaves["opacities"], np_(bg), 3, shs=leaves["shs"], scales=leaves["scales"], rotations=leaves["rotations"])
    assert np.array_equal(np_(out["radii"]), o["radii"])
    _images_close(out["render"], o["color"], "render")
    up = dict(c=torch.randn(3, H, W, device="cuda") / (3 * H * W), d=torch.randn(1, H, W, device="cuda") / (H * W),
              a=torch.randn(1, H, W, device="cuda") / (H * W))
    torch.autograd.backward((out["render"], out["depth"], out["alpha"]), (up["c"], up["d"], up["a"]))
    og = golden.backward(o, np_(up["c"]), np_(up["d"]), np_(up["a"]))
    for name, key in (("means3D", "dL_dmeans3D"), ("opacities", "dL_dopacity"), ("shs", "dL_dsh"), ("scales", "dL_dscales"),
                      ("rotations", "dL_drotations")):
        assert rel_inf(leaves[name].grad, og[key].reshape(leaves[name].grad.shape)) <= ORACLE_GRAD_RTOL, name
    assert rel_inf(out["viewspace_points"].grad, og["dL_dmeans2D"]) <= ORACLE_GRAD_RTOL
    # the other two switches of render(): Python-side SH evaluation and covariance, and an override colour
    pipe2 = types.SimpleNamespace(debug=False, compute_cov3D_python=True, convert_SHs_python=True)
    out2 = ref.gaussian_renderer.render(cam, pc, pipe2, bg)
    assert float((out2["render"] - out["render"]).abs().max()) < 2e-3
    flat = torch.full((5000, 3), 0.5, device="cuda")
    out3 = ref.gaussian_renderer.render(cam, pc, pipe, bg, override_color=flat)
    assert torch.equal(out3["radii"], out["radii"]) and torch.equal(out3["alpha"], out["alpha"])
