import ctypes, importlib, os, sys
import torch
sys.path.insert(0, os.getcwd())
importlib.import_module("3dgs_hierarchical_training_b200").install()
import htgs_native as nat
lib = nat.lib(); dev = torch.device("cuda:0")
for n, end_bit in ((1_000_000, 32), (765_027, 32), (3_383_305, 16), (3_383_305, 45), (3_383_305, 8)):
    g = torch.Generator().manual_seed(1)
    keys = torch.randint(0, 1 << min(end_bit, 62), (n,), generator=g, dtype=torch.int64).to(dev)
    vals = torch.arange(n, dtype=torch.int32, device=dev)
    ko, vo = torch.empty_like(keys), torch.empty_like(vals)
    scratch = torch.empty(lib.htgs_sort_scratch_bytes(n), dtype=torch.uint8, device=dev)
    st = ctypes.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
    kin, vin = keys.clone(), vals.clone()
    def run():
        nat.check(lib.htgs_sort_pairs(n, end_bit, kin.data_ptr(), vin.data_ptr(), ko.data_ptr(), vo.data_ptr(), scratch.data_ptr(), st), "sort")
    for _ in range(5): run()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(50): run()
    e1.record(); torch.cuda.synchronize()
    print(f"n={n} end_bit={end_bit}: {e0.elapsed_time(e1)/50*1e3:.1f} us")
