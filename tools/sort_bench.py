"""Times htgs_sort_pairs alone on rasterizer-shaped keys (tile<<32 | depth bits). Development aid."""
import ctypes, importlib, os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
importlib.import_module("3dgs_hierarchical_training_b200").install()
import htgs_native as nat

def main():
    lib = nat.lib()
    dev = torch.device("cuda:0")
    for n, end_bit in ((3_383_305, 45), (1_015_574, 43), (10_000_000, 45)):
        g = torch.Generator().manual_seed(1)
        tiles = 8160 if end_bit == 45 else 2040
        tile = torch.randint(0, tiles, (n,), generator=g, dtype=torch.int64)
        depth = (1.0 + 7.0 * torch.rand(n, generator=g)).float().view(torch.int32).long() & 0xFFFFFFFF
        keys = ((tile << 32) | depth).to(dev)
        vals = torch.arange(n, dtype=torch.int32, device=dev)
        ko, vo = torch.empty_like(keys), torch.empty_like(vals)
        scratch = torch.empty(lib.htgs_sort_scratch_bytes(n), dtype=torch.uint8, device=dev)
        st = ctypes.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
        kin, vin = keys.clone(), vals.clone()
        def run():
            kin.copy_(keys); vin.copy_(vals)
            nat.check(lib.htgs_sort_pairs(n, end_bit, kin.data_ptr(), vin.data_ptr(), ko.data_ptr(), vo.data_ptr(), scratch.data_ptr(), st), "sort")
        for _ in range(5): run()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        # time copies alone to subtract
        e0.record()
        for _ in range(20): kin.copy_(keys); vin.copy_(vals)
        e1.record(); torch.cuda.synchronize(); tc = e0.elapsed_time(e1) / 20
        e0.record()
        for _ in range(20): run()
        e1.record(); torch.cuda.synchronize(); t = e0.elapsed_time(e1) / 20 - tc
        ks, order = torch.sort(keys, stable=True)
        ok = torch.equal(ko, ks) and torch.equal(vo, order.int())
        e0.record()
        for _ in range(10): torch.sort(keys, stable=True)
        e1.record(); torch.cuda.synchronize(); tt = e0.elapsed_time(e1) / 10
        print(f"n={n} end_bit={end_bit} bits={os.environ.get('HTGS_SORT_DIGIT_BITS','9')}: {t*1e3:.1f} us  ({n*(8+24*((end_bit+7)//8))/t/1e6:.0f} GB/s alg)  ok={ok}  torch.sort(stable) {tt*1e3:.1f} us")

if __name__ == "__main__":
    main()
