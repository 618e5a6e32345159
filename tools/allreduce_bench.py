"""Times a 236 MB fp32 all-reduce (59 floats x 1M Gaussians) under torchrun. Development aid."""
import os, time, torch, torch.distributed as dist
rank = int(os.environ["RANK"]); world = int(os.environ["WORLD_SIZE"]); lr = int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(lr)
dist.init_process_group("nccl", device_id=torch.device("cuda", lr))
for n in (59_000_000, 59_000_000 * 6):
    x = torch.ones(n, device="cuda")
    for _ in range(5): dist.all_reduce(x)
    torch.cuda.synchronize(); dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(20): dist.all_reduce(x)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 20
    if rank == 0:
        print(f"world={world} bytes={4*n/1e6:.0f}MB allreduce {ms:.3f} ms algbw {4*n/ms/1e6:.0f} GB/s busbw {4*n/ms/1e6*2*(world-1)/world:.0f} GB/s  env ALGO={os.environ.get('NCCL_ALGO')} ", flush=True)
dist.destroy_process_group()
