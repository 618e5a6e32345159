"""GPU tests of the two stand-alone entry points: the onesweep radix sort (against torch's stable sort;
bit-exact, ragged sizes, partial key width) and distCUDA2 (bit-exact against the oracle / reference)."""
import ctypes
import os

import numpy as np
import pytest
import torch

import htgs_native as nat
import htgs_synthetic as syn
from oracle import golden, refcuda

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def _sort(keys, vals, end_bit):
    dev = keys.device
    n = keys.numel()
    lib = nat.lib()
    ko, vo = torch.empty_like(keys), torch.empty_like(vals)
    scratch = torch.empty(lib.htgs_sort_scratch_bytes(n), dtype=torch.uint8, device=dev)
    st = ctypes.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
    kin, vin = keys.clone(), vals.clone()  # the inputs are clobbered by the ping-pong passes
    nat.check(lib.htgs_sort_pairs(n, end_bit, kin.data_ptr(), vin.data_ptr(), ko.data_ptr(), vo.data_ptr(),
                                  scratch.data_ptr(), st), "htgs_sort_pairs")
    torch.cuda.synchronize()
    return ko, vo


@pytest.mark.parametrize("n", [1, 2, 31, 4095, 4096, 4097, 12289, 100_003, 3_000_000])
@pytest.mark.parametrize("end_bit", [41, 45, 30, 13, 64])
def test_onesweep_sort_matches_stable_sort(n, end_bit):
    g = torch.Generator().manual_seed(n * 131 + end_bit)
    hi = min(end_bit, 62)
    keys = torch.randint(0, 2 ** hi, (n,), generator=g, dtype=torch.int64)
    # many duplicate keys (stability) and a skewed digit distribution
    keys[::3] = keys[::3] & 0xFFFF
    keys[1::7] = keys[0]
    if end_bit < 64:
        keys = keys & ((1 << end_bit) - 1)
    keys = keys.cuda()
    vals = torch.arange(n, dtype=torch.int32, device="cuda")
    ko, vo = _sort(keys, vals, end_bit)
    ks, order = torch.sort(keys, stable=True)
    assert torch.equal(ko, ks)
    assert torch.equal(vo, order.int())


def test_sort_ignores_bits_above_end_bit():
    n, end_bit = 50_000, 20
    g = torch.Generator().manual_seed(5)
    keys = torch.randint(0, 2 ** 40, (n,), generator=g, dtype=torch.int64).cuda()
    vals = torch.arange(n, dtype=torch.int32, device="cuda")
    ko, vo = _sort(keys, vals, end_bit)
    low = keys & ((1 << end_bit) - 1)
    _, order = torch.sort(low, stable=True)
    assert torch.equal(vo, order.int()) and torch.equal(ko, keys[order])


@pytest.mark.parametrize("P", [1, 5, 700, 4097, 60_000])
def test_dist2_matches_oracle_bit_exact(P):
    from simple_knn._C import distCUDA2
    pts = syn.make_scene(max(P, 8), 64, 64, seed=90 + P)["means3D"][:P].contiguous()
    if P > 20_000:
        pts = pts[:20_000].contiguous()  # oracle is O(P^2)
    if pts.shape[0] >= 16:
        pts[5] = pts[9]  # exact duplicates count as distance 0 (self excluded by index)
    d = distCUDA2(pts.cuda())
    assert d.shape == (pts.shape[0],) and d.dtype == torch.float32 and d.is_cuda
    if pts.shape[0] >= 4:
        assert np.array_equal(d.cpu().numpy().view(np.uint32), golden.dist2(pts).view(np.uint32))


def test_dist2_large_against_reference_and_golden():
    from simple_knn._C import distCUDA2
    v = np.load(os.path.join(GOLD, "refcuda_knn.npz")) if os.path.exists(os.path.join(GOLD, "refcuda_knn.npz")) else None
    if v is not None:
        d = distCUDA2(torch.from_numpy(v["points"]).cuda())
        assert np.array_equal(d.cpu().numpy().view(np.uint32), v["dist2"].view(np.uint32))
    if refcuda.available():
        pts = syn.make_scene(300_000, 960, 540, seed=2)["means3D"].cuda()
        assert torch.equal(distCUDA2(pts), refcuda.dist2(pts))
