"""Merge-time operations (htgs_merge: calc_importance, importance_prune_mask, merge_two_models, capture / restore) against
the reference's OWN code run live: HTGaussianTrainer.calc_importance / merge_two_3DGS (trainer/ht3dgs_trainer.py:214-272,
:1428-1462) and HTGaussianModel.prune_points / densification_postfix / capture / restore (scene/gaussian_model_ht.py),
imported unchanged from baseline/_ref/htgs_callers (tools/stage_reference_callers.py) on top of this repository's plugin."""
import copy
import io
import math
import os
import sys
import types

import numpy as np
import pytest
import torch

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import ref_callers  # noqa: E402
from helpers import rel_inf  # noqa: E402

pytestmark = [pytest.mark.gpu,
              pytest.mark.skipif(not ref_callers.staged(), reason="baseline/_ref/htgs_callers not staged (tools/stage_reference_callers.py)")]

W, H = 160, 96
FOVX = 1.3541787529604106
OPT = types.SimpleNamespace(percent_dense=0.01, position_lr_init=1.6e-4, position_lr_final=1.6e-6, position_lr_delay_mult=0.01,
                            position_lr_max_steps=30000, feature_lr=2.5e-3, opacity_lr=0.05, scaling_lr=5e-3, rotation_lr=1e-3)


@pytest.fixture(scope="module")
def ref():
    ns = ref_callers.load()
    ns.trainer = ref_callers.load_trainer()
    return ns


def _camera(ref, uid=0, yaw=0.0):
    fovy = 2.0 * math.atan(math.tan(FOVX / 2.0) * H / W)
    c, s = math.cos(yaw), math.sin(yaw)
    R = np.array([[c, 0.0, s], [0.0, 1.0, 0.0], [-s, 0.0, c]], dtype=np.float64)
    T = np.array([0.02 * uid, -0.01, 0.2], dtype=np.float64)
    K = np.array([[W / (2 * math.tan(FOVX / 2)), 0, W / 2], [0, H / (2 * math.tan(fovy / 2)), H / 2], [0, 0, 1]], dtype=np.float64)
    return ref.cameras.Camera(colmap_id=uid, R=R, T=T, FoVx=FOVX, FoVy=fovy, image=torch.rand(3, H, W), gt_alpha_mask=None,
                              image_name=f"{uid:05d}", uid=uid, intrinsics=K, data_device="cuda")


def _render(ref, n=3000, seed=0, degree=3):
    g = np.random.default_rng(seed)
    z = g.uniform(1.5, 6.0, n)
    x = z * math.tan(FOVX / 2) * g.uniform(-1.1, 1.1, n)
    y = z * math.tan(FOVX / 2) * H / W * g.uniform(-1.1, 1.1, n)
    pcd = ref.graphics_utils.BasicPointCloud(points=np.stack([x, y, z], 1), colors=g.uniform(0, 1, (n, 3)), normals=np.zeros((n, 3)))
    r = ref.gaussian_model_ht.CF3DGS_Render(sh_degree=3, white_background=False, view_dependent=True)
    r.init_model(pcd)
    gm = r.gaussians
    torch.manual_seed(seed)
    gm._scaling.data += 0.9
    gm._rotation.data.normal_()
    gm._opacity.data.normal_(0.0, 1.5)
    gm._features_dc.data.normal_(0.0, 1.0)   # some colours leave [0, 1]: the clamp mask of the loss matters
    gm._features_rest.data.normal_(0.0, 0.25)
    gm.active_sh_degree = degree
    gm.training_setup(OPT)
    return r


@pytest.mark.parametrize("degree", [0, 2, 3])
def test_calc_importance_matches_reference(ref, degree):
    import htgs_merge
    r = _render(ref, degree=degree)
    cams = [_camera(ref, uid=i, yaw=0.04 * i - 0.04) for i in range(3)]
    pipe = types.SimpleNamespace(compute_cov3D_python=False, convert_SHs_python=False, debug=False)
    want = ref.trainer.HTGaussianTrainer.calc_importance(r, cams, pipe)  # autograd + abs() hooks through the plugin
    got = htgs_merge.calc_importance(r, cams, pipe)
    assert got.shape == want.shape == (3000, 48)
    assert float(want.max()) > 0
    assert rel_inf(got, want) <= 1e-5
    # element-wise: every entry within 1e-4 of its own value plus a floor of 1e-6 of the largest
    tol = 1e-4 * want.abs() + 1e-6 * want.abs().max()
    assert bool(((got - want).abs() <= tol).all())
    with pytest.raises(RuntimeError):
        htgs_merge.calc_importance(r, cams, types.SimpleNamespace(compute_cov3D_python=False, convert_SHs_python=True))


def test_importance_prune_mask_is_bottom_k():
    import htgs_merge
    g = torch.Generator().manual_seed(3)
    for P, ratio in ((10_000, 0.3), (4097, 0.5), (513, 0.01), (100, 0.0), (100, 1.0)):
        imp = torch.rand(P, 48, generator=g).cuda()
        imp[::7] = 0.0           # a block of exact ties at zero (culled Gaussians)
        imp[5::11] = imp[5]      # and ties at a positive value
        mask = htgs_merge.importance_prune_mask(imp, ratio)
        k = int(P * ratio)
        assert mask.dtype == torch.bool and int(mask.sum()) == k
        amax = imp.amax(-1)
        if 0 < k < P:
            assert float(amax[mask].max()) <= float(amax[~mask].min())
            # ties at the threshold go to the lower index (stable sort): identical to a stable argsort
            order = torch.sort(amax, stable=True).indices[:k]
            want = torch.zeros(P, dtype=torch.bool, device="cuda")
            want[order] = True
            assert torch.equal(mask, want)
        # without ties the choice is exactly torch.topk(largest=False), the reference's call
        imp2 = torch.rand(P, 48, generator=g).cuda() + 0.01
        m2 = htgs_merge.importance_prune_mask(imp2, ratio)
        want2 = torch.zeros(P, dtype=torch.bool, device="cuda")
        if k:
            want2[torch.topk(imp2.amax(-1), k, largest=False).indices] = True
        assert torch.equal(m2, want2)


def _train_steps(r, cam, n, fused):
    import htgs_merge
    if fused:
        htgs_merge.use_fused_adam(r.gaussians)
    for _ in range(n):
        pkg = r.render(cam)
        loss = (pkg["image"] - cam.original_image).abs().mean() + 0.01 * pkg["depth"].mean()
        loss.backward()
        r.gaussians.optimizer.step()
        r.gaussians.optimizer.zero_grad(set_to_none=True)


def test_merge_two_models_matches_reference_merge(ref):
    """merge_two_3DGS of the reference, run with a stand-in `self` (logger, camera loader, prune ratio) and its
    calc_importance replaced by a tie-free importance so that torch.topk's choice is unique, against
    htgs_merge.importance_prune_mask + merge_two_models on identical copies of the two models."""
    import htgs_merge
    Tr = ref.trainer.HTGaussianTrainer
    cam = _camera(ref)
    pairs = []
    for fused in (False, True):  # reference side keeps torch.optim.Adam; ours runs on FusedAdam
        dst, src = _render(ref, n=3000, seed=1), _render(ref, n=2500, seed=2)
        _train_steps(dst, cam, 2, fused)   # non-trivial Adam moments in the destination
        _train_steps(src, cam, 1, fused)
        pairs.append((dst, src))
    (dst_ref, src_ref), (dst_own, src_own) = pairs
    for a, b in ((dst_ref, dst_own), (src_ref, src_own)):
        for name in ("_xyz", "_features_dc", "_features_rest", "_opacity", "_scaling", "_rotation"):
            getattr(b.gaussians, name).data.copy_(getattr(a.gaussians, name).data)
        for ga, gb in zip(a.gaussians.optimizer.param_groups, b.gaussians.optimizer.param_groups):
            sa, sb = a.gaussians.optimizer.state[ga["params"][0]], b.gaussians.optimizer.state[gb["params"][0]]
            sb["exp_avg"].copy_(sa["exp_avg"])
            sb["exp_avg_sq"].copy_(sa["exp_avg_sq"])
    gen = torch.Generator().manual_seed(9)
    imp = {id(dst_ref): (torch.rand(3000, 48, generator=gen) + 0.01).cuda(), id(src_ref): (torch.rand(2500, 48, generator=gen) + 0.01).cuda()}
    T = torch.tensor([[0.98, -0.17, 0.05, 0.3], [0.17, 0.98, 0.02, -0.1], [-0.05, -0.01, 1.0, 0.2], [0.0, 0.0, 0.0, 1.0]], device="cuda")
    ratio = 0.3
    fake_self = types.SimpleNamespace(logger=types.SimpleNamespace(info=lambda *a, **k: None), pipe_cfg=types.SimpleNamespace(prune_ratio=ratio),
                                      load_viewpoint_cam=lambda fidx, pose=None, load_depth=True: cam)
    saved = Tr.calc_importance
    Tr.calc_importance = staticmethod(lambda gs_render, cameras, pipe: imp[id(gs_render)])
    try:
        Tr.merge_two_3DGS(fake_self, dst_ref, src_ref, T, to_visit_frames_dst=[0], to_visit_frames_src=[0])
    finally:
        Tr.calc_importance = saved
    mask_dst = htgs_merge.importance_prune_mask(imp[id(dst_ref)], ratio)
    mask_src = htgs_merge.importance_prune_mask(imp[id(src_ref)], ratio)
    added = htgs_merge.merge_two_models(dst_own.gaussians, src_own.gaussians, T, mask_dst, mask_src)
    a, b = dst_ref.gaussians, dst_own.gaussians
    n_dst = 3000 - int(3000 * ratio)
    assert added == 2500 - int(2500 * ratio) and b._xyz.shape[0] == a._xyz.shape[0] == n_dst + added
    assert torch.equal(a._xyz[:n_dst], b._xyz[:n_dst])                      # pruned destination rows: bit-exact
    assert rel_inf(b._xyz[n_dst:], a._xyz[n_dst:]) <= 1e-6                  # transformed source rows: matmul rounding only
    for name in ("_features_dc", "_features_rest", "_opacity", "_scaling", "_rotation"):
        assert torch.equal(getattr(a, name), getattr(b, name)), name
        assert isinstance(getattr(b, name), torch.nn.Parameter) and getattr(b, name).requires_grad
    for ga, gb in zip(a.optimizer.param_groups, b.optimizer.param_groups):
        assert ga["name"] == gb["name"] and gb["params"][0] is getattr(b, htgs_merge._ATTR[gb["name"]])
        sa, sb = a.optimizer.state[ga["params"][0]], b.optimizer.state[gb["params"][0]]
        assert torch.equal(sa["exp_avg"], sb["exp_avg"]) and torch.equal(sa["exp_avg_sq"], sb["exp_avg_sq"]), ga["name"]
        assert not sb["exp_avg"][n_dst:].any()
    for name in ("xyz_gradient_accum", "denom", "max_radii2D"):
        assert getattr(a, name).shape == getattr(b, name).shape and not getattr(b, name).any()
    # the merged model trains on
    _train_steps(dst_own, cam, 1, True)
    assert bool(torch.isfinite(dst_own.gaussians._xyz).all())


def test_capture_restore_round_trip_with_fused_adam(ref):
    """capture() -> torch.save -> torch.load -> restore() into a fresh model, then both continue training: identical bits.
    The tuple is also accepted by the reference's own restore (torch.optim.Adam) and vice versa."""
    import htgs_merge
    cam = _camera(ref)
    r = _render(ref, n=2000, seed=5)
    _train_steps(r, cam, 3, True)
    tup = htgs_merge.capture(r.gaussians)
    assert len(tup) == 13
    buf = io.BytesIO()
    torch.save(tup, buf)
    buf.seek(0)
    loaded = torch.load(buf, weights_only=False)
    r2 = ref.gaussian_model_ht.CF3DGS_Render(sh_degree=3, white_background=False, view_dependent=True)
    htgs_merge.restore(r2.gaussians, loaded, OPT)
    from htgs_optim import FusedAdam
    assert isinstance(r2.gaussians.optimizer, FusedAdam)
    assert r2.gaussians.active_sh_degree == r.gaussians.active_sh_degree
    st1, st2 = r.gaussians.optimizer.state_dict(), r2.gaussians.optimizer.state_dict()
    assert st1["param_groups"] == st2["param_groups"]
    for k in st1["state"]:
        assert torch.equal(st1["state"][k]["exp_avg"], st2["state"][k]["exp_avg"]) and st1["state"][k]["step"] == st2["state"][k]["step"]
    # the reference's own restore takes the same tuple
    r3 = ref.gaussian_model_ht.CF3DGS_Render(sh_degree=3, white_background=False, view_dependent=True)
    buf.seek(0)
    r3.gaussians.restore(torch.load(buf, weights_only=False), OPT)
    assert type(r3.gaussians.optimizer) is torch.optim.Adam
    assert torch.equal(r3.gaussians._xyz, r.gaussians._xyz)
    # continue: the restored model follows the original bit for bit (backward sums are order-nondeterministic, so compare
    # through a deterministic step: same gradients fed to both optimizers)
    for p1, p2 in zip((g["params"][0] for g in r.gaussians.optimizer.param_groups), (g["params"][0] for g in r2.gaussians.optimizer.param_groups)):
        grad = torch.randn_like(p1)
        p1.grad, p2.grad = grad.clone(), grad.clone()
    r.gaussians.optimizer.step()
    r2.gaussians.optimizer.step()
    for name in ("_xyz", "_features_dc", "_features_rest", "_opacity", "_scaling", "_rotation"):
        assert torch.equal(getattr(r.gaussians, name), getattr(r2.gaussians, name)), name
