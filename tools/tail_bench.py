"""Times the per-step tail at P = 1M Gaussians: this library's fused Adam and
densification-statistics kernels against what the reference runs (torch.optim.Adam default, masked
indexing) and against torch's own fused Adam.  Prints one JSON object."""
import importlib, json, os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
importlib.import_module("3dgs_hierarchical_training_b200").install()
import htgs_optim

SHAPES = dict(xyz=(3,), f_dc=(1, 3), f_rest=(15, 3), opacity=(1,), scaling=(3,), rotation=(4,))
LRS = dict(xyz=1.6e-4, f_dc=2.5e-3, f_rest=2.5e-3 / 20.0, opacity=5e-2, scaling=5e-3, rotation=1e-3)


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


def main(P=1_000_000):
    dev = "cuda"
    out = {"P": P}

    def make():
        ps = {k: torch.randn((P,) + s, device=dev).requires_grad_(True) for k, s in SHAPES.items()}
        for p in ps.values():
            p.grad = torch.randn_like(p)
        return ps, [{"params": [ps[k]], "lr": LRS[k], "name": k} for k in SHAPES]

    ps, groups = make()
    out["adam_torch_default_ms"] = timeit(torch.optim.Adam(groups, lr=0.0, eps=1e-15).step)
    ps, groups = make()
    out["adam_torch_fused_ms"] = timeit(torch.optim.Adam(groups, lr=0.0, eps=1e-15, fused=True).step)
    ps, groups = make()
    out["adam_htgs_fused_ms"] = timeit(htgs_optim.FusedAdam(groups, lr=0.0, eps=1e-15).step)
    nbytes = 59 * P * 4 * 7  # p, g, m, v read; p, m, v written
    out["adam_algorithmic_bytes"] = nbytes
    out["adam_htgs_GBps"] = nbytes / out["adam_htgs_fused_ms"] / 1e6

    radii = torch.randint(0, 30, (P,), device=dev, dtype=torch.int32) * (torch.rand(P, device=dev) > 0.25).int()
    grad = torch.randn(P, 3, device=dev)
    max_r, acc, den = torch.zeros(P, device=dev), torch.zeros(P, 1, device=dev), torch.zeros(P, 1, device=dev)

    def ref_stats():
        vis = radii > 0
        max_r[vis] = torch.max(max_r[vis], radii[vis])
        acc[vis] += torch.norm(grad[vis, :2], dim=-1, keepdim=True)
        den[vis] += 1

    out["stats_reference_expr_ms"] = timeit(ref_stats)
    out["stats_htgs_ms"] = timeit(lambda: htgs_optim.add_densification_stats(radii, grad, max_r, acc, den))
    print(json.dumps(out))


if __name__ == "__main__":
    main()
