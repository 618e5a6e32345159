"""fwd+bwd of one view from the model's STORED parameters: torch activations + cat feeding the default plugin call
(what scene/gaussian_model_ht.py:834-880 does) versus htgs_raw.rasterize_raw.  Prints one JSON line."""
import argparse
import json
import os
import sys

import torch

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="config3")
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--warmup", type=int, default=10)
    args = ap.parse_args()
    import importlib
    pkg = importlib.import_module("3dgs_hierarchical_training_b200")
    pkg.install()
    import diff_gaussian_rasterization as dgr
    import htgs_raw
    import htgs_synthetic as syn
    dev = torch.device("cuda", 0)
    cfg, scene, cam, up = syn.make_config(args.workload, device=dev)
    P, W, H = cfg["P"], cfg["W"], cfg["H"]
    leaves = dict(xyz=scene["means3D"].clone(), opacity=torch.logit(scene["opacities"].clamp(1e-6, 1 - 1e-6)),
                  features_dc=scene["shs"][:, :1, :].clone(), features_rest=scene["shs"][:, 1:, :].clone(),
                  scaling=torch.log(scene["scales"]), rotation=scene["rotations"].clone())
    leaves = {k: v.contiguous().requires_grad_(True) for k, v in leaves.items()}
    rs = dgr.GaussianRasterizationSettings(image_height=H, image_width=W, tanfovx=cam["tanfovx"], tanfovy=cam["tanfovy"],
                                           bg=torch.zeros(3, device=dev), scale_modifier=1.0, viewmatrix=cam["viewmatrix"],
                                           projmatrix=cam["projmatrix"], sh_degree=3, campos=cam["campos"], prefiltered=False,
                                           debug=False)
    ups = (up["dL_dcolor"], up["dL_ddepth"], up["dL_dalpha"])

    def composed():
        m2 = torch.zeros_like(leaves["xyz"], requires_grad=True)
        out = dgr.GaussianRasterizer(rs)(means3D=leaves["xyz"], means2D=m2, opacities=torch.sigmoid(leaves["opacity"]),
                                         shs=torch.cat((leaves["features_dc"], leaves["features_rest"]), dim=1),
                                         scales=torch.exp(leaves["scaling"]),
                                         rotations=torch.nn.functional.normalize(leaves["rotation"]))
        torch.autograd.backward((out[0], out[2], out[3]), ups)

    def raw():
        m2 = torch.zeros_like(leaves["xyz"], requires_grad=True)
        out = htgs_raw.rasterize_raw(leaves["xyz"], m2, leaves["opacity"], leaves["features_dc"], leaves["features_rest"],
                                     leaves["scaling"], leaves["rotation"], rs)
        torch.autograd.backward((out[0], out[2], out[3]), ups)

    res = {}
    for name, fn in (("torch_activations_plus_plugin", composed), ("raw_entry_points", raw)):
        for _ in range(args.warmup):
            for p in leaves.values():
                p.grad = None
            fn()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.reset_peak_memory_stats()
        e0.record()
        for _ in range(args.steps):
            for p in leaves.values():
                p.grad = None
            fn()
        e1.record()
        torch.cuda.synchronize()
        res[name] = {"ms_per_step": e0.elapsed_time(e1) / args.steps, "peak_alloc_MB": torch.cuda.max_memory_allocated() / 2**20}
    print(json.dumps({"workload": args.workload, "P": P, "W": W, "H": H, "steps": args.steps, **res}))


if __name__ == "__main__":
    main()
