"""GPU time of one full training step at config3 (render fwd, L1+SSIM, render bwd, densification statistics,
Adam): this repository's stack vs the reference stack (reference CUDA rasterizer from oracle/_ref + the
reference's torch loss formula + torch.optim.Adam + masked indexing).  Development/measurement aid; prints JSON."""
import importlib, json, os, sys
import torch
import torch.nn.functional as F
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
importlib.import_module("3dgs_hierarchical_training_b200").install()
import htgs_loss, htgs_optim, htgs_synthetic as syn
from diff_gaussian_rasterization import GaussianRasterizationSettings, GaussianRasterizer
from oracle import loss_ssim, refcuda


def timeit(fn, warm=5, it=30):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(it):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / it


def main(workload="config3"):
    dev = torch.device("cuda:0")
    cfg, scene, cam, _ = syn.make_config(workload, device=dev)
    H, W, P = cfg["H"], cfg["W"], cfg["P"]
    bg = torch.zeros(3, device=dev)
    target = torch.rand(3, H, W, device=dev)
    names = ("means3D", "shs", "opacities", "scales", "rotations")
    lrs = dict(means3D=1.6e-6, shs=2.5e-5, opacities=5e-4, scales=5e-5, rotations=1e-5)

    def fresh():
        p = {k: scene[k].clone().requires_grad_(True) for k in names}
        return p, [{"params": [p[k]], "lr": lrs[k], "name": k} for k in names]

    # ---- ours ----
    p, groups = fresh()
    opt = htgs_optim.FusedAdam(groups, lr=0.0, eps=1e-15)
    rs = GaussianRasterizationSettings(H, W, cam["tanfovx"], cam["tanfovy"], bg, 1.0, cam["viewmatrix"], cam["projmatrix"], 3,
                                       cam["campos"], False, False)
    rast = GaussianRasterizer(rs)
    max_r, acc, den = torch.zeros(P, device=dev), torch.zeros(P, 1, device=dev), torch.zeros(P, 1, device=dev)

    def ours():
        m2d = torch.zeros_like(p["means3D"], requires_grad=True)
        image, radii, depth, alpha = rast(means3D=p["means3D"], means2D=m2d, opacities=p["opacities"], shs=p["shs"],
                                          scales=p["scales"], rotations=p["rotations"])
        loss, _ = htgs_loss.fused_l1_ssim(image, target, 0.2)
        loss.backward()
        with torch.no_grad():
            htgs_optim.add_densification_stats(radii, m2d.grad, max_r, acc, den)
            opt.step()
            opt.zero_grad(set_to_none=True)

    out = {"workload": workload, "ours_ms": timeit(ours)}

    # ---- the same step from the model's STORED parameters (six optimizer groups, scene/gaussian_model_ht.py:275-287) ----
    import htgs_raw
    raw_lrs = dict(xyz=1.6e-6, features_dc=2.5e-5, features_rest=1.25e-6, opacity=5e-4, scaling=5e-5, rotation=1e-5)

    def fresh_raw():
        r = dict(xyz=scene["means3D"].clone(), features_dc=scene["shs"][:, :1].clone(), features_rest=scene["shs"][:, 1:].clone(),
                 opacity=torch.logit(scene["opacities"].clamp(1e-6, 1 - 1e-6)), scaling=torch.log(scene["scales"]),
                 rotation=scene["rotations"].clone())
        r = {k: v.contiguous().requires_grad_(True) for k, v in r.items()}
        return r, [{"params": [r[k]], "lr": raw_lrs[k], "name": k} for k in r]

    r, rgroups = fresh_raw()
    raw_opt = htgs_optim.FusedAdam(rgroups, lr=0.0, eps=1e-15)

    def ours_raw():
        m2d = torch.zeros_like(r["xyz"], requires_grad=True)
        image, radii, depth, alpha = htgs_raw.rasterize_raw(r["xyz"], m2d, r["opacity"], r["features_dc"], r["features_rest"],
                                                            r["scaling"], r["rotation"], rs)
        loss, _ = htgs_loss.fused_l1_ssim(image, target, 0.2)
        loss.backward()
        with torch.no_grad():
            htgs_optim.add_densification_stats(radii, m2d.grad, max_r, acc, den)
            raw_opt.step()
            raw_opt.zero_grad(set_to_none=True)

    out["ours_stored_params_ms"] = timeit(ours_raw)

    # ---- reference stack ----
    if refcuda.available():
        q, groups = fresh()
        ropt = torch.optim.Adam(groups, lr=0.0, eps=1e-15)
        w1 = torch.from_numpy(loss_ssim.window()).float().to(dev)
        win = (w1[:, None] @ w1[None, :]).expand(3, 1, 11, 11).contiguous()
        kw = dict(shs=q["shs"], scales=q["scales"], rotations=q["rotations"])

        def ref():
            fw = refcuda.forward(q["means3D"].detach(), q["opacities"].detach(), cam, bg, 3, want_state=False,
                                 **{k: v.detach() for k, v in kw.items()})
            image = fw["color"].clone().requires_grad_(True)
            x, y = image[None], target[None]
            mu1, mu2 = F.conv2d(x, win, padding=5, groups=3), F.conv2d(y, win, padding=5, groups=3)
            s1 = F.conv2d(x * x, win, padding=5, groups=3) - mu1.pow(2)
            s2 = F.conv2d(y * y, win, padding=5, groups=3) - mu2.pow(2)
            s12 = F.conv2d(x * y, win, padding=5, groups=3) - mu1 * mu2
            m = ((2 * mu1 * mu2 + 1e-4) * (2 * s12 + 9e-4)) / ((mu1.pow(2) + mu2.pow(2) + 1e-4) * (s1 + s2 + 9e-4))
            (0.8 * torch.abs(image - target).mean() + 0.2 * (1 - m.mean())).backward()
            zeros = torch.zeros(1, H, W, device=dev)
            g = refcuda.backward(fw, q["means3D"].detach(), cam, bg, image.grad, zeros, zeros, 3, **{k: v.detach() for k, v in kw.items()})
            q["means3D"].grad, q["shs"].grad, q["opacities"].grad = g["dL_dmeans3D"], g["dL_dsh"], g["dL_dopacity"]
            q["scales"].grad, q["rotations"].grad = g["dL_dscales"], g["dL_drotations"]
            with torch.no_grad():
                vis = fw["radii"] > 0
                max_r[vis] = torch.max(max_r[vis], fw["radii"][vis])
                acc[vis] += torch.norm(g["dL_dmeans2D"][vis, :2], dim=-1, keepdim=True)
                den[vis] += 1
                ropt.step()
                ropt.zero_grad(set_to_none=True)

        out["reference_stack_ms"] = timeit(ref)
        out["speedup"] = out["reference_stack_ms"] / out["ours_ms"]

        # reference stack from the stored parameters: torch activations + cat in front, autograd through them behind
        t, tgroups = fresh_raw()
        topt = torch.optim.Adam(tgroups, lr=0.0, eps=1e-15)

        def ref_raw():
            act = dict(shs=torch.cat((t["features_dc"], t["features_rest"]), dim=1), scales=torch.exp(t["scaling"]),
                       rotations=F.normalize(t["rotation"]))
            opac = torch.sigmoid(t["opacity"])
            det = {k: v.detach() for k, v in act.items()}
            fw = refcuda.forward(t["xyz"].detach(), opac.detach(), cam, bg, 3, want_state=False, **det)
            image = fw["color"].clone().requires_grad_(True)
            x, y = image[None], target[None]
            mu1, mu2 = F.conv2d(x, win, padding=5, groups=3), F.conv2d(y, win, padding=5, groups=3)
            s1 = F.conv2d(x * x, win, padding=5, groups=3) - mu1.pow(2)
            s2 = F.conv2d(y * y, win, padding=5, groups=3) - mu2.pow(2)
            s12 = F.conv2d(x * y, win, padding=5, groups=3) - mu1 * mu2
            m = ((2 * mu1 * mu2 + 1e-4) * (2 * s12 + 9e-4)) / ((mu1.pow(2) + mu2.pow(2) + 1e-4) * (s1 + s2 + 9e-4))
            (0.8 * torch.abs(image - target).mean() + 0.2 * (1 - m.mean())).backward()
            zeros = torch.zeros(1, H, W, device=dev)
            g = refcuda.backward(fw, t["xyz"].detach(), cam, bg, image.grad, zeros, zeros, 3, **det)
            t["xyz"].grad = g["dL_dmeans3D"]
            torch.autograd.backward((act["shs"], act["scales"], act["rotations"], opac),
                                    (g["dL_dsh"], g["dL_dscales"], g["dL_drotations"], g["dL_dopacity"]))
            with torch.no_grad():
                vis = fw["radii"] > 0
                max_r[vis] = torch.max(max_r[vis], fw["radii"][vis])
                acc[vis] += torch.norm(g["dL_dmeans2D"][vis, :2], dim=-1, keepdim=True)
                den[vis] += 1
                topt.step()
                topt.zero_grad(set_to_none=True)

        out["reference_stack_stored_params_ms"] = timeit(ref_raw)
        out["speedup_stored_params"] = out["reference_stack_stored_params_ms"] / out["ours_stored_params_ms"]
    print(json.dumps(out))


if __name__ == "__main__":
    main(sys.argv[1] if len(sys.argv) > 1 else "config3")
