"""Margin of this library's gradients against the element-wise bar of tests/helpers.py, per tensor, several runs.

  python tools/grad_margin.py [workload ...]        (GPU + oracle/_ref; test infrastructure, not product)
Prints one JSON line per workload: {tensor: [elem_excess of run 1, run 2, run 3]} against ONE run of the unmodified
reference CUDA -- the figure the parity test asserts to be <= 1."""
import importlib
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
importlib.import_module("3dgs_hierarchical_training_b200").install()

import torch  # noqa: E402

import htgs_synthetic as syn  # noqa: E402
from helpers import GRAD_NAMES, elem_excess, rel_inf, run_ours  # noqa: E402
from oracle import refcuda  # noqa: E402


def main():
    dev = torch.device("cuda:0")
    for w in (sys.argv[1:] or ["config2", "config3", "config5"]):
        cfg, scene, cam, up = syn.make_config(w, device=dev)
        bg = torch.tensor([0.1, 0.2, 0.3], device=dev)
        kw = dict(shs=scene["shs"], scales=scene["scales"], rotations=scene["rotations"])
        r = refcuda.forward(scene["means3D"], scene["opacities"], cam, bg, 3, **kw)
        rg = refcuda.backward(r, scene["means3D"], cam, bg, up["dL_dcolor"], up["dL_ddepth"], up["dL_dalpha"], 3, **kw)
        out = {k: [] for k in GRAD_NAMES if k in rg}
        worst_rel = 0.0
        for _ in range(3):
            mine = run_ours(scene, cam, up, bg)
            for k in out:
                g = mine["grads"][k]
                out[k].append(round(elem_excess(g, rg[k].view(g.shape)), 3))
                worst_rel = max(worst_rel, rel_inf(g, rg[k].view(g.shape)))
        # the worst element of the worst tensor of the last run: who it is and how large the disagreement is
        worst = max(out, key=lambda k: out[k][-1])
        g, ref = mine["grads"][worst].reshape(-1).double(), rg[worst].reshape(-1).double()
        scale = float(ref.abs().max())
        ex = (g - ref).abs() / (1e-3 * ref.abs() + 1e-5 * scale)
        i = int(ex.argmax())
        per = g.numel() // scene["means3D"].shape[0]
        gid = i // per
        detail = dict(tensor=worst, gaussian=gid, component=i % per, ours=float(g[i]), ref=float(ref[i]), tensor_max=scale,
                      abs_err_over_max=float((g[i] - ref[i]).abs() / scale), rel_err=float((g[i] - ref[i]).abs() / ref[i].abs().clamp_min(1e-30)),
                      elements_over_half_bar=int((ex > 0.5).sum()), radius=int(mine["radii"][gid]),
                      mean2D=[float(v) for v in mine["means2D"][gid]], conic_opacity=[float(v) for v in mine["conic_opacity"][gid]],
                      row_ours=[float(v) for v in mine["grads"][worst].reshape(-1, per)[gid]],
                      row_ref=[float(v) for v in rg[worst].reshape(-1, per)[gid]])
        refcuda.dgr().ref_release()
        print(json.dumps({"workload": w, "elem_excess": out, "worst_rel_inf": worst_rel, "worst_element": detail}))


if __name__ == "__main__":
    main()
