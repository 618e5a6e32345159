"""Run under torchrun on N GPUs: checks htgs_allreduce_p2p (FlatGradReducer(peer_memory=True)) against NCCL's
all-reduce on the same data and times both.  Rank 0 prints one JSON line."""
import importlib, json, os, sys
import torch
import torch.distributed as dist
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
importlib.import_module("3dgs_hierarchical_training_b200").install()
import htgs_dist


def main():
    rank, world, lr = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    dev = torch.device("cuda", lr)
    torch.cuda.set_device(dev)
    dist.init_process_group("nccl", device_id=dev)
    P = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
    mode = sys.argv[2] if len(sys.argv) > 2 else True  # "p2p" | "multimem" | default (automatic)
    red = htgs_dist.FlatGradReducer(P, dev, peer_memory=mode)
    g = torch.Generator(device=dev).manual_seed(10 + rank)
    data = torch.randn(red.flat.numel(), generator=g, device=dev)
    ref = data.clone()
    dist.all_reduce(ref)
    red.flat.copy_(data)
    red.all_reduce()
    torch.cuda.synchronize()
    err = float((red.flat - ref).abs().max() / ref.abs().max())
    # replicas identical?
    chk = red.flat.double().sum().reshape(1)
    lst = [torch.zeros_like(chk) for _ in range(world)]
    dist.all_gather(lst, chk)
    identical = all(float(x) == float(lst[0]) for x in lst)

    def timeit(fn, it=30):
        for _ in range(5):
            fn()
        dist.barrier(); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(it):
            fn()
        e1.record(); torch.cuda.synchronize()
        t = torch.tensor([e0.elapsed_time(e1) / it], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t)

    t_p2p = timeit(red.all_reduce)
    t_nccl = timeit(lambda: dist.all_reduce(ref))
    if rank == 0:
        nbytes = red.flat.numel() * 4
        print(json.dumps({"collective": red.collective, "world": world, "bytes": nbytes, "rel_err_vs_nccl": err, "replicas_identical": identical,
                          "p2p_ms": t_p2p, "nccl_ms": t_nccl, "p2p_busbw_GBps": 2 * (world - 1) / world * nbytes / t_p2p / 1e6,
                          "nccl_busbw_GBps": 2 * (world - 1) / world * nbytes / t_nccl / 1e6}))
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
