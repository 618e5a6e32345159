#!/usr/bin/env python
"""Gradient self-noise of the reference CUDA rasterizer and the distance of this library's gradients from it.

The reference's backward sums per-pixel contributions with float atomics in whatever order the hardware
schedules them, so two runs of the SAME reference binary on the SAME inputs differ.  This tool runs the
unmodified reference CUDA (oracle/_ref) twice and this library once on each named workload and records, per
gradient tensor,
  * rel_inf      = ||a - b||_inf / ||b||_inf                      (the 1e-4 bar of BASELINE.json)
  * elem(r, a)   = max_i |a_i - b_i| / (r * |b_i| + a * ||b||_inf) (element-wise: small entries are held to their own
                   magnitude plus an absolute floor tied to the tensor's scale)
for (ref run 2 vs ref run 1) and (ours vs ref run 1).  Output: one JSON document on stdout / --out.
Needs a GPU and oracle/_ref; test infrastructure, not product."""
import argparse
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
from helpers import run_ours  # noqa: E402
from oracle import refcuda  # noqa: E402

NAMES = ["dL_dmeans2D", "dL_dopacity", "dL_dmeans3D", "dL_dsh", "dL_dscales", "dL_drotations"]
COMBOS = [(1e-3, 1e-6), (1e-2, 1e-6), (1e-2, 1e-5), (1e-3, 1e-5), (1e-4, 1e-5)]


def measures(a, b):
    a, b = a.reshape(-1).double(), b.reshape(-1).double()
    d = (a - b).abs()
    scale = float(b.abs().max().clamp_min(1e-30))
    out = {"rel_inf": float(d.max()) / scale, "scale": scale}
    for r, at in COMBOS:
        out[f"elem_r{r:g}_a{at:g}"] = float((d / (r * b.abs() + at * scale)).max())
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workloads", default="config2,config3,config5")
    ap.add_argument("--out", default=None)
    args = ap.parse_args()
    dev = torch.device("cuda:0")
    doc = {"what": "reference CUDA gradient self-noise (two runs of oracle/_ref) and ours vs reference", "workloads": {}}
    for name in args.workloads.split(","):
        cfg, scene, cam, up = syn.make_config(name, device=dev)
        bg = torch.zeros(3, device=dev)
        kw = dict(shs=scene["shs"], scales=scene["scales"], rotations=scene["rotations"])
        refs = []
        for _ in range(2):
            fw = refcuda.forward(scene["means3D"], scene["opacities"], cam, bg, 3, want_state=False, **kw)
            g = refcuda.backward(fw, scene["means3D"], cam, bg, up["dL_dcolor"], up["dL_ddepth"], up["dL_dalpha"], 3, **kw)
            torch.cuda.synchronize()
            refs.append({k: g[k].clone() for k in NAMES})
        mine = run_ours(scene, cam, up, bg)["grads"]
        mine2 = run_ours(scene, cam, up, bg)["grads"]
        ent = {}
        for k in NAMES:
            ent[k] = {"ref_vs_ref": measures(refs[1][k], refs[0][k]), "ours_vs_ref": measures(mine[k].view_as(refs[0][k]), refs[0][k]),
                      "ours_vs_ours": measures(mine2[k], mine[k])}
        doc["workloads"][name] = ent
        del scene, up, refs, mine, mine2
        torch.cuda.empty_cache()
    txt = json.dumps(doc, indent=1)
    if args.out:
        with open(args.out, "w") as f:
            f.write(txt)
    print(txt)


if __name__ == "__main__":
    main()
