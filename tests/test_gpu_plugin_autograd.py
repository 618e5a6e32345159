"""The autograd contract of the drop-in plugin: 4-tuple, gradient slots,
retain_grad on the dummy means2D, in-place edits of the returned depth/colour, zero-materialised
grads for unused outputs, accumulation over several views (what view-parallel training relies on)."""
import numpy as np
import pytest
import torch

import htgs_synthetic as syn
from helpers import GRAD_RTOL, np_, rel_inf, run_ours
import htgs_native as nat
from oracle import golden

pytestmark = pytest.mark.gpu


def _setup(cfgname="small"):
    from diff_gaussian_rasterization import GaussianRasterizationSettings, GaussianRasterizer
    dev = torch.device("cuda:0")
    cfg, scene, cam, up = syn.make_config(cfgname, device=dev)
    bg = torch.zeros(3, device=dev)
    rs = GaussianRasterizationSettings(image_height=cfg["H"], image_width=cfg["W"], tanfovx=cam["tanfovx"], tanfovy=cam["tanfovy"],
                                       bg=bg, scale_modifier=1.0, viewmatrix=cam["viewmatrix"], projmatrix=cam["projmatrix"],
                                       sh_degree=3, campos=cam["campos"], prefiltered=False, debug=False)
    return cfg, scene, cam, up, bg, GaussianRasterizer(raster_settings=rs)


def test_render_like_cf3dgs_render():
    """Mimics CF3DGS_Render.render (scene/gaussian_model_ht.py:775-908) and train_step's use of it."""
    cfg, scene, cam, up, bg, rast = _setup()
    xyz = scene["means3D"].clone().requires_grad_(True)
    raw_op = torch.logit(scene["opacities"].clamp(1e-4, 1 - 1e-4)).requires_grad_(True)
    raw_sc = torch.log(scene["scales"]).requires_grad_(True)
    raw_rot = (scene["rotations"] * 1.7).requires_grad_(True)
    f_dc = scene["shs"][:, :1].clone().requires_grad_(True)
    f_rest = scene["shs"][:, 1:].clone().requires_grad_(True)
    screenspace_points = torch.zeros_like(xyz, dtype=xyz.dtype, requires_grad=True, device="cuda") + 0
    screenspace_points.retain_grad()
    out = rast(means3D=xyz, means2D=screenspace_points, shs=torch.cat((f_dc, f_rest), dim=1), colors_precomp=None,
               opacities=torch.sigmoid(raw_op), scales=torch.exp(raw_sc), rotations=torch.nn.functional.normalize(raw_rot),
               cov3D_precomp=None)
    assert len(out) == 4
    rendered_image, radii, depth, alpha = out
    assert rendered_image.shape == (3, cfg["H"], cfg["W"]) and depth.shape == alpha.shape == (1, cfg["H"], cfg["W"])
    assert radii.dtype == torch.int32 and not radii.requires_grad
    image = rendered_image.clamp(0, 1)
    depth[depth < 0.5] = 0.5  # in-place edit of the plugin's output (trainer/ht3dgs_trainer.py:1292)
    loss = (image * up["dL_dcolor"]).sum() + (depth * up["dL_ddepth"]).sum()  # alpha unused -> zero grad materialised
    loss.backward()
    for t in (xyz, raw_op, raw_sc, raw_rot, f_dc, f_rest):
        assert t.grad is not None and bool(torch.isfinite(t.grad).all())
    vsp = screenspace_points.grad
    assert vsp.shape == xyz.shape and not vsp[:, 2].any()          # z slot never written
    vis = radii > 0
    assert not vsp[~vis].any() and bool(vsp[vis].abs().sum() > 0)  # densification stat source
    assert bool(torch.norm(vsp[vis, :2], dim=-1).isfinite().all())


def test_gradient_accumulates_over_views_and_matches_oracle():
    """Two cameras, one backward each, parameter grads accumulate -- the per-GPU half of view-parallel training."""
    from diff_gaussian_rasterization import GaussianRasterizationSettings, GaussianRasterizer
    dev = torch.device("cuda:0")
    cfg, scene, _, up = syn.make_config("small", device=dev)
    bg = torch.tensor([0.1, 0.3, 0.2], device=dev)
    params = {k: v.clone().requires_grad_(True) for k, v in scene.items()}
    total = {k: 0.0 for k in ("dL_dmeans3D", "dL_dsh", "dL_dopacity", "dL_dscales", "dL_drotations")}
    for yaw in (-4.0, 5.0):
        cam = syn.make_camera(cfg["W"], cfg["H"], yaw_deg=yaw, device=dev)
        rs = GaussianRasterizationSettings(cfg["H"], cfg["W"], cam["tanfovx"], cam["tanfovy"], bg, 1.0, cam["viewmatrix"],
                                           cam["projmatrix"], 3, cam["campos"], False, False)
        color, radii, depth, alpha = GaussianRasterizer(rs)(
            means3D=params["means3D"], means2D=torch.zeros_like(params["means3D"], requires_grad=True),
            opacities=params["opacities"], shs=params["shs"], scales=params["scales"], rotations=params["rotations"])
        torch.autograd.backward((color, depth, alpha), (up["dL_dcolor"], up["dL_ddepth"], up["dL_dalpha"]))
        o = golden.forward(scene["means3D"], scene["opacities"], cam, bg, 3, shs=scene["shs"], scales=scene["scales"],
                           rotations=scene["rotations"])
        og = golden.backward(o, up["dL_dcolor"], up["dL_ddepth"], up["dL_dalpha"])
        for k in total:
            total[k] = total[k] + og[k].astype(np.float64)
    got = dict(dL_dmeans3D=params["means3D"].grad, dL_dsh=params["shs"].grad, dL_dopacity=params["opacities"].grad,
               dL_dscales=params["scales"].grad, dL_drotations=params["rotations"].grad)
    for k, v in got.items():
        assert rel_inf(v, total[k].reshape(np_(v).shape)) <= 5e-3, k


def test_forward_only_no_grad_and_colors_precomp_grad_slot():
    cfg, scene, cam, up, bg, rast = _setup("tiny")
    with torch.no_grad():
        out = rast(means3D=scene["means3D"], means2D=torch.zeros_like(scene["means3D"]), opacities=scene["opacities"],
                   shs=scene["shs"], scales=scene["scales"], rotations=scene["rotations"])
    assert not out[0].requires_grad
    colors = torch.rand(cfg["P"], 3, device="cuda").requires_grad_(True)
    color, radii, depth, alpha = rast(means3D=scene["means3D"], means2D=torch.zeros_like(scene["means3D"]),
                                      opacities=scene["opacities"], colors_precomp=colors, scales=scene["scales"],
                                      rotations=scene["rotations"])
    color.sum().backward()
    assert colors.grad is not None and colors.grad.shape == colors.shape and bool((colors.grad >= 0).all())
    assert not colors.grad[radii == 0].any()


def test_gradient_arena_is_adopted_and_equal():
    """htgs_dist.FlatGradReducer.install_arena: the backward writes into the flat all-reduce buffer and autograd
    adopts those views as .grad (no pack copy); values equal the ordinary path."""
    import htgs_dist
    cfg, scene, cam, up, bg, rast = _setup("small")
    dev = torch.device("cuda:0")
    P = cfg["P"]

    def run():
        p = {k: scene[k].clone().requires_grad_(True) for k in htgs_dist.PARAM_ORDER}
        out = rast(means3D=p["means3D"], means2D=torch.zeros_like(p["means3D"], requires_grad=True), opacities=p["opacities"],
                   shs=p["shs"], scales=p["scales"], rotations=p["rotations"])
        torch.autograd.backward((out[0], out[2], out[3]), (up["dL_dcolor"], up["dL_ddepth"], up["dL_dalpha"]))
        return p

    plain = run()
    red = htgs_dist.FlatGradReducer(P, dev)
    red.install_arena()
    try:
        arena = run()
        copied = red.pack_if_needed({k: arena[k].grad for k in htgs_dist.PARAM_ORDER})
    finally:
        red.remove_arena()
    assert copied == 0, "autograd did not adopt the arena views"
    views = red.unpack({k: plain[k] for k in htgs_dist.PARAM_ORDER})
    for k in htgs_dist.PARAM_ORDER:
        assert rel_inf(views[k], plain[k].grad) <= GRAD_RTOL, k
    again = run()  # arena removed: ordinary allocation again
    assert again["shs"].grad.data_ptr() != red.flat.data_ptr() + 4 * red.slices["shs"][0]


def test_gradient_arena_with_live_grad_accumulates_instead_of_overwriting():
    """A parameter that still holds last step's arena slice as .grad (zero_grad(set_to_none=False), or two views per
    step) must end with the SUM of both backwards, as plain autograd gives -- not with a gradient added to itself."""
    import htgs_dist
    cfg, scene, cam, up, bg, rast = _setup("small")
    dev = torch.device("cuda:0")
    p = {k: scene[k].clone().requires_grad_(True) for k in htgs_dist.PARAM_ORDER}

    def backward_once(scale):
        out = rast(means3D=p["means3D"], means2D=torch.zeros_like(p["means3D"], requires_grad=True), opacities=p["opacities"],
                   shs=p["shs"], scales=p["scales"], rotations=p["rotations"])
        torch.autograd.backward((out[0], out[2], out[3]), (scale * up["dL_dcolor"], scale * up["dL_ddepth"], scale * up["dL_dalpha"]))

    backward_once(1.0)
    single = {k: p[k].grad.clone() for k in p}
    for v in p.values():
        v.grad = None
    red = htgs_dist.FlatGradReducer(cfg["P"], dev)
    red.install_arena(p)
    try:
        backward_once(1.0)   # adopted: .grad is the arena slice
        assert p["shs"].grad.data_ptr() == red.flat.data_ptr() + 4 * red.slices["shs"][0]
        backward_once(2.0)   # .grad still live: must accumulate 1x + 2x into the same slice
        assert red.pack_if_needed({k: p[k].grad for k in htgs_dist.PARAM_ORDER}) == 0
    finally:
        red.remove_arena()
    for k in htgs_dist.PARAM_ORDER:
        assert rel_inf(p[k].grad, 3.0 * single[k]) <= 2 * GRAD_RTOL, k


def test_speculative_forward_reruns_when_the_predicted_capacity_is_too_small():
    """diff_gaussian_rasterization/_C.py sizes the binning buffer from the previous call's R.  Poisoning that
    prediction far too low forces the overflow branch (phase 2 redone at the exact size); while the undersized
    attempt runs, the kernels must stay inside the small buffer (guards in make_keys / sort / tile ranges).  The
    result must be bit-identical to the blocking two-phase flow, forward and backward state included."""
    from diff_gaussian_rasterization import _C
    cfg, scene, cam, up = syn.make_config("config1", device="cuda")
    bg = torch.tensor([0.1, 0.2, 0.3], device="cuda")
    P, W, H = cfg["P"], cfg["W"], cfg["H"]
    key = (0, P, W, H)
    saved = _C.SPECULATIVE
    try:
        _C.SPECULATIVE = False
        exact = run_ours(scene, cam, up, bg)
        _C.SPECULATIVE = True
        R = exact["R"]
        assert R > 16000  # so that capacity = 1.25 * poison + 8192 really is too small below
        for poison in (R // 8, 1, R):   # far too small, absurdly small, exact (no rerun)
            _C._R_PREDICTION[key] = poison
            # canary behind the undersized buffer would be torch's next block; instead check every output bit
            got = run_ours(scene, cam, up, bg)
            assert got["R"] == R and _C._R_PREDICTION[key] == R
            for k in ("color", "depth", "alpha", "radii", "n_contrib", "ranges", "tiles_touched", "point_offsets"):
                assert torch.equal(got[k], exact[k]), (poison, k)
            assert torch.equal(got["keys"][:R], exact["keys"][:R]) and torch.equal(got["point_list"][:R], exact["point_list"][:R])
            for k in ("dL_dmeans3D", "dL_dsh", "dL_dopacity", "dL_dscales", "dL_drotations", "dL_dmeans2D"):
                assert rel_inf(got["grads"][k], exact["grads"][k]) <= GRAD_RTOL, (poison, k)
    finally:
        _C.SPECULATIVE = saved
        _C._R_PREDICTION.pop(key, None)


def test_backward_reuses_cached_reach_masks_only_under_the_forwards_carving():
    """The forward blend caches the per-instance reach masks in the binning buffer; the backward may use them only when
    it carves that buffer with the count the forward used (`num_rendered.carve`).  Handing it the exact R after a
    capacity-sized (speculative) forward must silently fall back to recomputing them: same gradients either way."""
    from diff_gaussian_rasterization import _C
    cfg, scene, cam, up = syn.make_config("config1", device="cuda")
    bg = torch.tensor([0.1, 0.2, 0.3], device="cuda")
    empty = torch.Tensor([])
    args = (bg, scene["means3D"], empty, scene["opacities"], scene["scales"], scene["rotations"], 1.0, empty, cam["viewmatrix"],
            cam["projmatrix"], cam["tanfovx"], cam["tanfovy"], cfg["H"], cfg["W"], scene["shs"], 3, cam["campos"], False, False)
    _C.rasterize_gaussians(*args)            # first call: exact two-phase flow, seeds the prediction
    R, color, depth, alpha, radii, geom, binning, img = _C.rasterize_gaussians(*args)   # speculative: capacity-sized buffer
    assert isinstance(R, int) and R.carve > int(R)

    def bwd(count):
        g = _C.rasterize_gaussians_backward(bg, scene["means3D"], radii, empty, scene["scales"], scene["rotations"], 1.0, empty,
                                            cam["viewmatrix"], cam["projmatrix"], cam["tanfovx"], cam["tanfovy"], up["dL_dcolor"],
                                            up["dL_ddepth"], up["dL_dalpha"], scene["shs"], 3, cam["campos"], geom, count, binning, img,
                                            alpha, False)
        torch.cuda.synchronize()
        return g

    cached, recomputed = bwd(R), bwd(int(R))
    for a, b in zip(cached, recomputed):
        assert rel_inf(a, b) <= GRAD_RTOL
    # and a cache written under another carving is not trusted: poison it, the fallback must not read it
    lay = nat.binning_layout(int(R), cfg["W"], cfg["H"])
    binning[lay["values_alt"] + 4 * int(R):lay["sort_scratch"]] = 0   # where the exact-R carving would look for masks
    for a, b in zip(bwd(int(R)), cached):
        assert rel_inf(a, b) <= GRAD_RTOL
