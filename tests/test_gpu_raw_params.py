"""Raw-parameter entry points (htgs_forward_preprocess_raw_async / htgs_backward_raw)
against the composition they replace: torch activations + torch.cat feeding the parity-tested default plugin path,
with autograd walking back through those torch ops (scene/gaussian_model_ht.py:127-188, :834-866)."""
import numpy as np
import pytest
import torch

import htgs_synthetic as syn
from helpers import IMAGE_ATOL, GRAD_RTOL, np_, rel_inf

pytestmark = pytest.mark.gpu


def _raw_leaves(scene, seed):
    g = torch.Generator().manual_seed(seed)
    P = scene["means3D"].shape[0]
    stretch = (0.25 + 3.0 * torch.rand(P, 1, generator=g)).to(scene["rotations"].device)  # stored quaternions are not unit
    leaves = dict(xyz=scene["means3D"].clone(), opacity=torch.logit(scene["opacities"].clamp(1e-6, 1 - 1e-6)),
                  features_dc=scene["shs"][:, :1, :].clone(), features_rest=scene["shs"][:, 1:, :].clone(),
                  scaling=torch.log(scene["scales"]), rotation=scene["rotations"] * stretch)
    return {k: v.contiguous().requires_grad_(True) for k, v in leaves.items()}


def _settings(dgr, cfg, cam, bg, deg):
    return dgr.GaussianRasterizationSettings(image_height=cfg["H"], image_width=cfg["W"], tanfovx=cam["tanfovx"],
                                             tanfovy=cam["tanfovy"], bg=bg, scale_modifier=1.0, viewmatrix=cam["viewmatrix"],
                                             projmatrix=cam["projmatrix"], sh_degree=deg, campos=cam["campos"],
                                             prefiltered=False, debug=False)


@pytest.mark.parametrize("cfgname,deg", [("tiny", 3), ("small", 3), ("small", 0), ("small", 1), ("small", 2), ("ragged", 3),
                                         ("config1", 3)])
def test_raw_entry_points_match_activated_composition(cfgname, deg):
    import diff_gaussian_rasterization as dgr
    import htgs_raw
    dev = torch.device("cuda", 0)
    cfg, scene, cam, up = syn.make_config(cfgname, device=dev)
    bg = torch.tensor([0.3, 0.6, 0.9], device=dev)
    rs = _settings(dgr, cfg, cam, bg, deg)
    ups = (up["dL_dcolor"], up["dL_ddepth"], up["dL_dalpha"])

    a = _raw_leaves(scene, 77)
    m2a = torch.zeros_like(a["xyz"], requires_grad=True)
    out_a = dgr.GaussianRasterizer(rs)(means3D=a["xyz"], means2D=m2a, opacities=torch.sigmoid(a["opacity"]),
                                       shs=torch.cat((a["features_dc"], a["features_rest"]), dim=1),
                                       scales=torch.exp(a["scaling"]), rotations=torch.nn.functional.normalize(a["rotation"]))
    torch.autograd.backward((out_a[0], out_a[2], out_a[3]), ups)

    b = _raw_leaves(scene, 77)
    m2b = torch.zeros_like(b["xyz"], requires_grad=True)
    out_b = htgs_raw.rasterize_raw(b["xyz"], m2b, b["opacity"], b["features_dc"], b["features_rest"], b["scaling"],
                                   b["rotation"], rs)
    torch.autograd.backward((out_b[0], out_b[2], out_b[3]), ups)

    # the activations are evaluated with the same fp32 formulas torch uses, up to the summation order of the
    # quaternion norm; a last-bit difference there may move a radius by one in rare cases
    P = cfg["P"]
    assert int((out_a[1] != out_b[1]).sum()) <= max(0, int(1e-4 * P))
    same_radii = bool((out_a[1] == out_b[1]).all())
    for i, name in ((0, "color"), (2, "depth"), (3, "alpha")):
        d = (out_a[i] - out_b[i]).abs().max().item()
        tol = IMAGE_ATOL * max(1.0, out_a[i].abs().max().item())
        assert d <= tol or not same_radii, f"{name}: {d}"
    tol = GRAD_RTOL if same_radii else 5e-3
    assert rel_inf(m2b.grad, m2a.grad) <= tol
    for k in a:
        assert b[k].grad is not None and b[k].grad.shape == a[k].grad.shape, k
        assert rel_inf(b[k].grad, a[k].grad) <= tol, k


def test_raw_rest_rows_are_written_exactly_once_and_nowhere_else():
    """The 180-byte _features_rest rows are moved with aligned vector accesses for the interior and scalar
    accesses at the edges; every phase (row index mod 4) and both ends of the array are covered here."""
    import diff_gaussian_rasterization as dgr
    import htgs_raw
    dev = torch.device("cuda", 0)
    cfg, scene, cam, up = syn.make_config("small", device=dev)
    rs = _settings(dgr, cfg, cam, torch.zeros(3, device=dev), 3)
    for P in (1, 2, 3, 5, 1030):
        sub = {k: v[:P].contiguous() for k, v in scene.items()}
        b = _raw_leaves(sub, 3)
        m2 = torch.zeros_like(b["xyz"], requires_grad=True)
        out = htgs_raw.rasterize_raw(b["xyz"], m2, b["opacity"], b["features_dc"], b["features_rest"], b["scaling"],
                                     b["rotation"], rs)
        torch.autograd.backward((out[0], out[2], out[3]), (up["dL_dcolor"], up["dL_ddepth"], up["dL_dalpha"]))
        a = _raw_leaves(sub, 3)
        m2a = torch.zeros_like(a["xyz"], requires_grad=True)
        oa = dgr.GaussianRasterizer(rs)(means3D=a["xyz"], means2D=m2a, opacities=torch.sigmoid(a["opacity"]),
                                        shs=torch.cat((a["features_dc"], a["features_rest"]), dim=1),
                                        scales=torch.exp(a["scaling"]), rotations=torch.nn.functional.normalize(a["rotation"]))
        torch.autograd.backward((oa[0], oa[2], oa[3]), (up["dL_dcolor"], up["dL_ddepth"], up["dL_dalpha"]))
        assert torch.isfinite(b["features_rest"].grad).all()
        assert rel_inf(b["features_rest"].grad, a["features_rest"].grad) <= GRAD_RTOL or a["features_rest"].grad.abs().max() == 0
        assert rel_inf(b["features_dc"].grad, a["features_dc"].grad) <= GRAD_RTOL or a["features_dc"].grad.abs().max() == 0


def test_raw_rejects_host_tensors_and_bad_shapes():
    import diff_gaussian_rasterization as dgr
    import htgs_raw
    dev = torch.device("cuda", 0)
    cfg, scene, cam, up = syn.make_config("tiny", device=dev)
    rs = _settings(dgr, cfg, cam, torch.zeros(3, device=dev), 3)
    b = {k: v.detach() for k, v in _raw_leaves(scene, 1).items()}
    m2 = torch.zeros_like(b["xyz"])
    with pytest.raises(RuntimeError):
        htgs_raw.rasterize_raw(b["xyz"].cpu(), m2, b["opacity"], b["features_dc"], b["features_rest"], b["scaling"], b["rotation"], rs)
    with pytest.raises(RuntimeError):
        htgs_raw.rasterize_raw(b["xyz"], m2, b["opacity"], b["features_dc"], b["features_rest"][:, :8], b["scaling"], b["rotation"], rs)
    with pytest.raises(RuntimeError):
        htgs_raw.rasterize_raw(b["xyz"], m2, b["opacity"][:-1], b["features_dc"], b["features_rest"], b["scaling"], b["rotation"], rs)
