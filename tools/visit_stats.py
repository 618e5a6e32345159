"""How many (warp, Gaussian) visits would finer pixel blocks save?  Design study for the blend kernels, run on the box.

For a sample of tiles of a workload: the exact per-pixel liveness of every listed Gaussian (alpha >= 1/255, power <= 0,
list position below the pixel's last contributor), reduced to pixel blocks of several shapes.  Prints, per block shape,
  visits        = sum over blocks of Gaussians with at least one live pixel in the block,
  lanes/visit   = live pixels per visit,
and for a 32-lane warp made of k independent sub-blocks that walk their own lists side by side,
  paired visits = sum over warps of max over its sub-blocks of the sub-block's list length.

  python tools/visit_stats.py [workload] [n_tiles]
"""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "3dgs_hierarchical_training_b200"))
sys.path.insert(0, os.path.join(ROOT, "tests"))
sys.path.insert(0, ROOT)

import htgs_synthetic as syn  # noqa: E402
from helpers import run_ours  # noqa: E402


def main():
    workload = sys.argv[1] if len(sys.argv) > 1 else "config3"
    n_tiles = int(sys.argv[2]) if len(sys.argv) > 2 else 600
    dev = torch.device("cuda:0")
    cfg, scene, cam, up = syn.make_config(workload, device=dev)
    bg = torch.zeros(3, device=dev)
    out = run_ours(scene, cam, up, bg)
    W, H = cam["W"], cam["H"]
    gx, gy = (W + 15) // 16, (H + 15) // 16
    g = torch.Generator().manual_seed(0)
    tiles = torch.randperm(gx * gy, generator=g)[:n_tiles].tolist()
    ys, xs = torch.meshgrid(torch.arange(16, device=dev), torch.arange(16, device=dev), indexing="ij")
    # block shapes (h, w) and how many of them make a 32-lane warp
    shapes = {"8x4 (now)": (4, 8), "8x2": (2, 8), "4x4": (4, 4), "4x2": (2, 4), "2x2": (2, 2)}
    visits = {k: 0 for k in shapes}
    live_total = 0
    paired = {"2 x 8x2 in an 8x4 warp": 0, "2 x 4x4 in an 8x4 warp": 0, "4 x 4x2 in an 8x4 warp": 0, "8 x 2x2 in an 8x4 warp": 0}
    listed = 0
    for t in tiles:
        r0, r1 = (int(v) for v in out["ranges"][t])
        if r1 <= r0:
            continue
        ids = out["point_list"][r0:r1].long()
        n = ids.numel()
        listed += n
        tx, ty = t % gx, t // gx
        px = (tx * 16 + xs).float()[None]
        py = (ty * 16 + ys).float()[None]
        m = out["means2D"][ids]
        co = out["conic_opacity"][ids]
        dx = m[:, 0, None, None] - px
        dy = m[:, 1, None, None] - py
        power = -0.5 * (co[:, 0, None, None] * dx * dx + co[:, 2, None, None] * dy * dy) - co[:, 1, None, None] * dx * dy
        alpha = torch.clamp(co[:, 3, None, None] * torch.exp(power), max=0.99)
        inside = ((tx * 16 + xs) < W) & ((ty * 16 + ys) < H)
        ncon = torch.zeros(16, 16, dtype=torch.long, device=dev)
        yy, xx = (ty * 16 + ys).clamp(max=H - 1), (tx * 16 + xs).clamp(max=W - 1)
        ncon = out["n_contrib"][yy, xx].long() * inside
        pos = torch.arange(n, device=dev)[:, None, None]
        live = (power <= 0) & (alpha >= 1.0 / 255.0) & (pos < ncon[None])  # [n, 16, 16]
        live_total += int(live.sum())
        per = {}
        for name, (h, w) in shapes.items():
            b = live.view(n, 16 // h, h, 16 // w, w).any(dim=4).any(dim=2)  # [n, by, bx]
            per[name] = b
            visits[name] += int(b.sum())

        def pair(name, sub, fh, fw):  # sub-blocks of shape `sub` grouped fh x fw into one 8x4 warp block
            b = per[sub]  # [n, by, bx]
            cnt = b.sum(dim=0)  # list length per sub-block
            by, bx = cnt.shape
            c = cnt.view(by // fh, fh, bx // fw, fw)
            paired[name] += int(c.amax(dim=3).amax(dim=1).sum())

        pair("2 x 8x2 in an 8x4 warp", "8x2", 2, 1)
        pair("2 x 4x4 in an 8x4 warp", "4x4", 1, 2)
        pair("4 x 4x2 in an 8x4 warp", "4x2", 2, 2)
        pair("8 x 2x2 in an 8x4 warp", "2x2", 2, 4)
    res = dict(workload=workload, tiles_sampled=len(tiles), listed_instances=listed, live_pairs=live_total,
               visits={k: v for k, v in visits.items()},
               live_lanes_per_visit={k: round(live_total / max(v, 1), 2) for k, v in visits.items()},
               warp_visits_with_independent_sub_blocks=paired,
               relative_to_now={k: round(v / max(visits["8x4 (now)"], 1), 3) for k, v in paired.items()})
    print(json.dumps(res))


if __name__ == "__main__":
    main()
