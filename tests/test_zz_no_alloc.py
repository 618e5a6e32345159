"""htgs_sort_pairs_static: the pair sort for callers that may not allocate device memory.

Some users of this library live inside a memory budget that is not theirs to extend -- a plugin
inside a host application's pre-reserved arena, a serving process whose pool is fixed and accounted,
a latency-critical loop where cudaMalloc would stall the device.  They can hand over inputs and
outputs they already own but cannot provide (or obtain) scratch, so this entry point keeps its
scratch in a fixed array inside the library and therefore caps the input at
HTGS_SORT_STATIC_MAX_PAIRS.

The test calls it through the C ABI and demands the same bytes as the ordinary entry point
(htgs_sort_pairs with caller-provided scratch), twice in a row with different data, because the
second call inherits whatever the first left in the shared array.  It also checks that an oversized
request is refused with a message rather than run.  This file sorts last on purpose: a run that
stops at the first failure has then already executed every other test."""
import ctypes

import pytest
import torch

import htgs_native as nat

pytestmark = pytest.mark.gpu


def _stream(dev):
    return ctypes.c_void_p(torch.cuda.current_stream(dev).cuda_stream)


def _reference_sort(keys, vals, end_bit):
    """The existing entry point: scratch allocated by the caller."""
    lib = nat.lib()
    n = keys.numel()
    kin, vin = keys.clone(), vals.clone()
    ko, vo = torch.empty_like(keys), torch.empty_like(vals)
    scratch = torch.empty(lib.htgs_sort_scratch_bytes(n), dtype=torch.uint8, device=keys.device)
    nat.check(lib.htgs_sort_pairs(n, end_bit, kin.data_ptr(), vin.data_ptr(), ko.data_ptr(), vo.data_ptr(), scratch.data_ptr(),
                                  _stream(keys.device)), "htgs_sort_pairs")
    torch.cuda.synchronize()
    return ko, vo


def _static_sort(keys, vals, end_bit):
    """The no-allocation entry point: only inputs and outputs cross the boundary."""
    lib = nat.lib()
    n = keys.numel()
    kin, vin = keys.clone(), vals.clone()
    ko, vo = torch.empty_like(keys), torch.empty_like(vals)
    nat.check(lib.htgs_sort_pairs_static(n, end_bit, kin.data_ptr(), vin.data_ptr(), ko.data_ptr(), vo.data_ptr(),
                                         _stream(keys.device)), "htgs_sort_pairs_static")
    torch.cuda.synchronize()
    return ko, vo


def _data(n, end_bit, seed):
    g = torch.Generator().manual_seed(seed)
    keys = torch.randint(0, 2 ** min(end_bit, 62), (n,), generator=g, dtype=torch.int64)
    keys[::5] = keys[0]  # duplicates: the order of equal keys must match too (stable sort)
    return keys.cuda(), torch.arange(n, dtype=torch.int32, device="cuda")


def test_static_scratch_sort_equals_allocating_sort_twice_in_a_row():
    # first call: as many pairs as the built-in scratch allows, 45 key bits -> 6 radix passes (result ends in the input
    # arrays and is copied out); second call: ragged size, 33 bits -> 5 passes (result lands in the outputs directly)
    for n, end_bit, seed in ((nat.SORT_STATIC_MAX_PAIRS, 45, 1), (300_007, 33, 2)):
        keys, vals = _data(n, end_bit, seed)
        ko_ref, vo_ref = _reference_sort(keys, vals, end_bit)
        ko, vo = _static_sort(keys, vals, end_bit)
        assert torch.equal(ko, ko_ref), f"keys differ (n={n}, end_bit={end_bit})"
        assert torch.equal(vo, vo_ref), f"values differ (n={n}, end_bit={end_bit})"
        ks, order = torch.sort(keys, stable=True)
        assert torch.equal(ko, ks) and torch.equal(vo, order.int())


def test_static_scratch_sort_leaves_device_free_memory_unchanged():
    """The property the entry point exists for: a call takes no device memory, not even transiently.
    The driver's own free-memory figure (cudaMemGetInfo, which sees every allocation on the device, torch's
    included) must read the same before and after.  One call is made first and not measured: it is the call
    that resolves the array's address, and with CUDA's lazy module loading it is also the moment the library
    image -- the 2 MB array included -- is placed on the device."""
    lib = nat.lib()
    dev = torch.device("cuda", torch.cuda.current_device())
    n, end_bit = 200_003, 45
    keys, vals = _data(n, end_bit, 7)
    _static_sort(keys, vals, end_bit)  # first use on this device (unmeasured, see above)
    kin, vin = keys.clone(), vals.clone()
    ko, vo = torch.empty_like(keys), torch.empty_like(vals)
    torch.cuda.synchronize()
    free_before, _ = torch.cuda.mem_get_info(dev)
    reserved_before = torch.cuda.memory_reserved(dev)
    for _ in range(3):
        nat.check(lib.htgs_sort_pairs_static(n, end_bit, kin.data_ptr(), vin.data_ptr(), ko.data_ptr(), vo.data_ptr(),
                                             _stream(dev)), "htgs_sort_pairs_static")
        free_during, _ = torch.cuda.mem_get_info(dev)  # GPU work still in flight
        assert free_during == free_before
        kin.copy_(keys)
        vin.copy_(vals)
    torch.cuda.synchronize()
    free_after, _ = torch.cuda.mem_get_info(dev)
    assert free_after == free_before, f"device free memory moved by {free_before - free_after} bytes"
    assert torch.cuda.memory_reserved(dev) == reserved_before  # and torch's caching allocator was not touched either
    ks, order = torch.sort(keys, stable=True)
    assert torch.equal(ko, ks) and torch.equal(vo, order.int())


def test_static_scratch_sort_refuses_more_than_its_maximum():
    lib = nat.lib()
    n = nat.SORT_STATIC_MAX_PAIRS + 1
    keys, vals = _data(n, 40, 3)
    ko, vo = torch.empty_like(keys), torch.empty_like(vals)
    rc = lib.htgs_sort_pairs_static(n, 40, keys.data_ptr(), vals.data_ptr(), ko.data_ptr(), vo.data_ptr(), _stream(keys.device))
    assert rc != 0
    msg = lib.htgs_last_error().decode()
    assert "HTGS_SORT_STATIC_MAX_PAIRS" in msg and str(n) in msg
    # and the array is still usable afterwards
    keys, vals = _data(4097, 40, 4)
    ko_ref, vo_ref = _reference_sort(keys, vals, 40)
    ko, vo = _static_sort(keys, vals, 40)
    assert torch.equal(ko, ko_ref) and torch.equal(vo, vo_ref)
