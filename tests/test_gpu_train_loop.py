"""End-to-end: a miniature HTGaussianTrainer.train_step loop (render -> loss -> backward -> densification
statistics -> Adam) built only from this repository's CUDA pieces, fitting a perturbed scene back to a target
image.  Checks that the pieces compose (shapes, dtypes, gradient flow through the activations the model applies,
scene/gaussian_model_ht.py:826-839) and that optimisation actually reduces the loss."""
import pytest
import torch

import htgs_synthetic as syn

pytestmark = pytest.mark.gpu


def test_training_loop_reduces_loss():
    import htgs_loss
    import htgs_optim
    from diff_gaussian_rasterization import GaussianRasterizationSettings, GaussianRasterizer
    dev = torch.device("cuda:0")
    cfg, scene, cam, _ = syn.make_config("small", device=dev)
    H, W, P = cfg["H"], cfg["W"], cfg["P"]
    bg = torch.zeros(3, device=dev)
    rs = GaussianRasterizationSettings(H, W, cam["tanfovx"], cam["tanfovy"], bg, 1.0, cam["viewmatrix"], cam["projmatrix"], 3,
                                       cam["campos"], False, False)
    rast = GaussianRasterizer(rs)

    def render(p):
        means2D = torch.zeros_like(p["xyz"], requires_grad=True)
        means2D.retain_grad()
        out = rast(means3D=p["xyz"], means2D=means2D, opacities=torch.sigmoid(p["opacity"]),
                   shs=torch.cat((p["f_dc"], p["f_rest"]), dim=1), scales=torch.exp(p["scaling"]),
                   rotations=torch.nn.functional.normalize(p["rotation"]))
        return out, means2D

    truth = dict(xyz=scene["means3D"], f_dc=scene["shs"][:, :1].contiguous(), f_rest=scene["shs"][:, 1:].contiguous(),
                 opacity=torch.logit(scene["opacities"].clamp(1e-3, 1 - 1e-3)), scaling=torch.log(scene["scales"]),
                 rotation=scene["rotations"])
    with torch.no_grad():
        (target, _, _, _), _ = render(truth)
        target = target.clamp(0, 1)
    g = torch.Generator().manual_seed(3)
    params = {k: (v + 0.15 * torch.randn(v.shape, generator=g).to(dev) * (0.3 if k == "xyz" else 1.0)).detach().requires_grad_(True)
              for k, v in truth.items()}
    lrs = dict(xyz=2e-3, f_dc=1e-2, f_rest=5e-4, opacity=5e-2, scaling=5e-3, rotation=1e-3)
    opt = htgs_optim.FusedAdam([{"params": [params[k]], "lr": lrs[k], "name": k} for k in params], lr=0.0, eps=1e-15)
    max_radii2D = torch.zeros(P, device=dev)
    grad_accum, denom = torch.zeros(P, 1, device=dev), torch.zeros(P, 1, device=dev)
    losses = []
    for it in range(40):
        (image, radii, depth, alpha), means2D = render(params)
        loss, parts = htgs_loss.fused_l1_ssim(image, target, 0.2)
        loss.backward()
        with torch.no_grad():
            htgs_optim.add_densification_stats(radii, means2D.grad, max_radii2D, grad_accum, denom)
            opt.step()
            opt.zero_grad(set_to_none=True)
        losses.append(float(loss.detach()))
    assert all(torch.isfinite(p).all() for p in params.values())
    assert losses[-1] < 0.6 * losses[0], losses[::8]
    assert min(losses[-5:]) <= min(losses[:5])
    vis_any = denom.squeeze(1) > 0
    assert bool(vis_any.any()) and bool((max_radii2D[vis_any] > 0).all())
    assert float(denom.max()) <= 40 and bool((grad_accum[~vis_any] == 0).all())
