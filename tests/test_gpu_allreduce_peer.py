"""Own all-reduce kernels (htgs_allreduce_p2p / htgs_allreduce_multimem) against NCCL on 2, 4 and 8 GPUs of one node.
Skipped on a single-GPU box; the host-side logic of the reducer is covered by the gloo tests."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

pytestmark = pytest.mark.gpu


def _worker(rank, world, port, mode, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world), LOCAL_RANK=str(rank))
    import importlib
    importlib.import_module("3dgs_hierarchical_training_b200").install()
    import htgs_dist
    dev = torch.device("cuda", rank)
    torch.cuda.set_device(dev)
    dist.init_process_group("nccl", device_id=dev)
    try:
        P = 10_007  # 59 * P is not a multiple of 4: exercises the padded tail
        red = htgs_dist.FlatGradReducer(P, dev, peer_memory=mode)
        g = torch.Generator(device=dev).manual_seed(3 + rank)
        ok = True
        for it in range(3):  # repeated use of the same buffer and barriers
            data = torch.randn(red.flat.numel(), generator=g, device=dev)
            ref = data.clone()
            dist.all_reduce(ref)
            red.flat.copy_(data)
            red.all_reduce()
            torch.cuda.synchronize()
            if world == 2:
                ok = ok and torch.equal(red.flat, ref)  # two ranks: a + b is the same in any order
            else:  # the switch (multimem) and NCCL choose their own summation order from 3 ranks up
                ok = ok and float((red.flat - ref).abs().max() / ref.abs().max()) <= 1e-6
            # every replica must hold the same bits whatever the order was
            mine = red.flat.clone()
            lo, hi = mine.clone(), mine.clone()
            dist.all_reduce(lo, op=dist.ReduceOp.MIN)
            dist.all_reduce(hi, op=dist.ReduceOp.MAX)
            ok = ok and torch.equal(lo, hi)
        out.put((rank, ok, red.collective))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 4, 8])
@pytest.mark.parametrize("mode", ["p2p", "multimem"])
def test_peer_memory_allreduce_matches_nccl(mode, world):
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    ctx = mp.get_context("spawn")
    out = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, world, port, mode, out)) for r in range(world)]
    for p in procs:
        p.start()
    res = [out.get(timeout=300) for _ in range(world)]
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    assert all(ok for _, ok, _ in res)
    assert all(c == mode for _, _, c in res)
