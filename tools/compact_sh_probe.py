"""Development aid: times the two kernels the compact SH exchange adds (htgs_sh_expand for 1..8 view slots, and the
backward in compact mode against the ordinary backward) on one GPU at config 3."""
import ctypes, importlib, json, os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
importlib.import_module("3dgs_hierarchical_training_b200").install()
import htgs_native as nat, htgs_synthetic as syn, htgs_dist
from helpers import run_ours

def ev(fn, n=20):
    for _ in range(3): fn()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); e0.record()
    for _ in range(n): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n

def main():
    dev = torch.device("cuda:0")
    cfg, scene, cam, up = syn.make_config("config3", device=dev)
    P = cfg["P"]; lib = nat.lib(); out = {}
    st = ctypes.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
    for V in (1, 2, 4, 8):
        red = htgs_dist.CompactSHGradReducer(P, dev, views_total=V)
        red.buf[red.n_small_pad:].normal_()
        slots = (ctypes.c_void_p * V)(*[red.buf.data_ptr() + 4 * red._slot_offset(v) for v in range(V)])
        out[f"expand_ms_views{V}"] = ev(lambda: nat.check(lib.htgs_sh_expand(P, 3, 16, V, scene["means3D"].data_ptr(), slots, red.sh_grad.data_ptr(), st), "x"))
    bg = torch.zeros(3, device=dev)
    out["fwd_bwd_plain_ms"] = ev(lambda: run_ours(scene, cam, up, bg), 10)
    red = htgs_dist.CompactSHGradReducer(P, dev, views_total=1)
    red.install(None)
    try:
        def step():
            red.begin_step(); run_ours(scene, cam, up, bg)
        out["fwd_bwd_compact_ms"] = ev(step, 10)
    finally:
        red.remove()
    print(json.dumps(out))

if __name__ == "__main__":
    main()
