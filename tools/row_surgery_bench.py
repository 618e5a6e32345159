"""prune_points / densify_and_clone tensor work at P = 1M: reference torch indexing vs the fused row kernels."""
import importlib, json, os, sys, time
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
importlib.import_module("3dgs_hierarchical_training_b200").install()
import htgs_optim
import test_gpu_row_surgery as T


def wall(fn, it=5):
    fn(); torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(it):
        fn()
    torch.cuda.synchronize()
    return 1e3 * (time.perf_counter() - t0) / it


def main(P=1_000_000):
    out = {"P": P}
    for op, frac in (("prune", 0.9), ("clone", 0.1)):
        mask = (torch.rand(P, generator=torch.Generator().manual_seed(3)) < frac).cuda()
        stats = [torch.rand(P, 1).cuda(), torch.rand(P, 1).cuda(), torch.rand(P).cuda()]

        def ref():
            o = T._make_optimizer(P, htgs_optim.FusedAdam)
            torch.cuda.synchronize(); t0 = time.perf_counter()
            if op == "prune":
                T._reference_prune(o, mask); [s[mask] for s in stats]
            else:
                T._reference_clone(o, mask)
            torch.cuda.synchronize(); return 1e3 * (time.perf_counter() - t0)

        def mine():
            o = T._make_optimizer(P, htgs_optim.FusedAdam)
            torch.cuda.synchronize(); t0 = time.perf_counter()
            if op == "prune":
                htgs_optim.prune_optimizer(o, mask, extra=stats)
            else:
                htgs_optim.clone_selected_to_optimizer(o, mask)
            torch.cuda.synchronize(); return 1e3 * (time.perf_counter() - t0)

        ref(); mine()
        out[op] = {"torch_indexing_ms": min(ref() for _ in range(5)), "htgs_rows_ms": min(mine() for _ in range(5))}
    print(json.dumps(out))


if __name__ == "__main__":
    main()
