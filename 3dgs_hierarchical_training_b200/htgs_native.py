"""ctypes binding of libhtgs.so (include/htgs.h).

There is no fallback: if the CUDA library is missing or a call fails this
module raises.  Tensors are allocated by the caller with torch; only raw
pointers and sizes cross the boundary.
"""
import ctypes
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libhtgs.so")

c_void_p, c_int, c_int64, c_float, c_size_t = ctypes.c_void_p, ctypes.c_int, ctypes.c_int64, ctypes.c_float, ctypes.c_size_t

class AdamGroup(ctypes.Structure):
    """htgs_adam_group of include/htgs.h."""
    _fields_ = [("param", c_void_p), ("grad", c_void_p), ("exp_avg", c_void_p), ("exp_avg_sq", c_void_p),
                ("numel", ctypes.c_longlong), ("lr", c_float)]


class RowTensor(ctypes.Structure):
    """htgs_row_tensor of include/htgs.h."""
    _fields_ = [("src", c_void_p), ("dst", c_void_p), ("row_words", c_int), ("zero_new", c_int)]


ROWS_MAX_TENSORS = 24  # HTGS_ROWS_MAX_TENSORS

_SIGNATURES = {
    "htgs_abi_version": (c_int, []),
    "htgs_last_error": (ctypes.c_char_p, []),
    "htgs_geom_bytes": (c_size_t, [c_int]),
    "htgs_image_bytes": (c_size_t, [c_int, c_int]),
    "htgs_binning_bytes": (c_size_t, [c_int64, c_int, c_int]),
    "htgs_backward_scratch_bytes": (c_size_t, [c_int]),
    "htgs_geom_layout": (c_int, [c_int, ctypes.POINTER(c_size_t), c_int]),
    "htgs_binning_layout": (c_int, [c_int64, c_int, c_int, ctypes.POINTER(c_size_t), c_int]),
    "htgs_image_layout": (c_int, [c_int, c_int, ctypes.POINTER(c_size_t), c_int]),
    "htgs_forward_preprocess": (c_int, [c_int, c_int, c_int, c_void_p, c_void_p, c_float, c_void_p, c_void_p, c_void_p,
                                        c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_float, c_float,
                                        c_int, c_void_p, c_void_p, ctypes.POINTER(c_int64), c_void_p]),
    "htgs_forward_preprocess_async": (c_int, [c_int, c_int, c_int, c_void_p, c_void_p, c_float, c_void_p, c_void_p, c_void_p,
                                              c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_float, c_float,
                                              c_int, c_void_p, c_void_p, c_void_p, c_void_p]),
    "htgs_forward_render_capacity": (c_int, [c_int, c_int64, c_int, c_int, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p,
                                             c_void_p, c_void_p, c_void_p, c_void_p]),
    "htgs_forward_render": (c_int, [c_int, c_int64, c_int, c_int, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p,
                                    c_void_p, c_void_p, c_void_p, c_void_p]),
    "htgs_backward": (c_int, [c_int, c_int, c_int, c_int64, c_int, c_int, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p,
                              c_void_p, c_float, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_float, c_float,
                              c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p,
                              c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p]),
    # (P, D, xyz, scaling_raw, scale_modifier, rotation_raw, opacity_raw, features_dc, features_rest, view, proj, campos,
    #  W, H, tanfovx, tanfovy, prefiltered, geom, radii, pinned_count, stream)
    "htgs_forward_preprocess_raw_async": (c_int, [c_int, c_int, c_void_p, c_void_p, c_float, c_void_p, c_void_p, c_void_p,
                                                  c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_float, c_float,
                                                  c_int, c_void_p, c_void_p, c_void_p, c_void_p]),
    # (P, D, R, W, H, bg, xyz, f_dc, f_rest, alphas, scaling_raw, scale_modifier, rotation_raw, opacity_raw, view, proj,
    #  campos, tanfovx, tanfovy, radii, geom, binning, image, dL_dcolor, dL_ddepth, dL_dalpha, scratch, 7 gradient outputs, stream)
    "htgs_backward_raw": (c_int, [c_int, c_int, c_int64, c_int, c_int, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p,
                                  c_void_p, c_float, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_float, c_float,
                                  c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p,
                                  c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p]),
    "htgs_mark_visible": (c_int, [c_int, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p]),
    "htgs_allreduce_p2p": (c_int, [ctypes.POINTER(c_void_p), c_int, c_int, ctypes.c_longlong, c_void_p]),
    "htgs_allreduce_multimem": (c_int, [c_void_p, c_int, c_int, ctypes.c_longlong, c_void_p]),
    "htgs_allreduce_set_launch_shape": (c_int, [c_int, c_int]),
    "htgs_sh_slot_floats": (c_size_t, [c_int]),
    # (P, D, M, R, W, H, bg, means3D, shs, alphas, scales, scale_modifier, rotations, cov3D, view, proj, campos, tanfovx, tanfovy,
    #  radii, geom, binning, image, dL_dcolor, dL_ddepth, dL_dalpha, scratch, dmeans2D, dopacity, dmeans3D, dcov3D, dscales, drot,
    #  rgb_slot, is_multicast, stream)
    "htgs_backward_compact": (c_int, [c_int, c_int, c_int, c_int64, c_int, c_int, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p,
                                      c_float, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_float, c_float, c_void_p, c_void_p,
                                      c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p,
                                      c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_void_p]),
    "htgs_sh_expand": (c_int, [c_int, c_int, c_int, c_int, c_void_p, ctypes.POINTER(c_void_p), c_void_p, c_void_p]),
    # (P, D, M, R, W, H, bg, means3D, campos, alphas, radii, geom, binning, image, dL_dcolor, scratch, importance, stream)
    "htgs_importance_accumulate": (c_int, [c_int, c_int, c_int, c_int64, c_int, c_int, c_void_p, c_void_p, c_void_p, c_void_p,
                                           c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p]),
    "htgs_importance_select_workspace_bytes": (c_size_t, [c_int]),
    "htgs_importance_select": (c_int, [c_int, c_int, c_void_p, c_int64, c_void_p, c_void_p, c_void_p]),
    "htgs_transform_points": (c_int, [c_int, c_void_p, c_void_p, c_void_p, c_void_p]),
    "htgs_row_scan_scratch_bytes": (c_size_t, [c_int]),
    "htgs_row_scan": (c_int, [c_int, c_void_p, c_void_p, c_void_p, c_void_p]),
    "htgs_rows_compact": (c_int, [c_int, c_void_p, c_void_p, c_int, ctypes.POINTER(RowTensor), c_void_p]),
    "htgs_rows_append": (c_int, [c_int, c_void_p, c_void_p, c_int, ctypes.POINTER(RowTensor), c_void_p]),
    "htgs_rows_split": (c_int, [c_int, c_int, ctypes.c_uint32, c_void_p, c_void_p, c_int, ctypes.POINTER(RowTensor), c_void_p]),
    "htgs_sort_scratch_bytes": (c_size_t, [c_int64]),
    "htgs_sort_pairs": (c_int, [c_int64, c_int, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p]),
    "htgs_sort_pairs_static": (c_int, [c_int64, c_int, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p]),
    "htgs_profile_enable": (None, [c_int]),
    "htgs_profile_read": (c_int, [ctypes.POINTER(ctypes.c_double), ctypes.POINTER(c_int), c_int]),
    "htgs_blend_stats": (c_int, [c_int, c_int64, c_int, c_int, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p]),
    "htgs_densification_stats": (c_int, [c_int, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p]),
    "htgs_adam_step": (c_int, [c_int, ctypes.POINTER(AdamGroup), c_float, c_float, c_float, ctypes.c_longlong, c_void_p]),
    "htgs_l1_ssim_scratch_bytes": (c_size_t, [c_int, c_int, c_int]),
    "htgs_l1_ssim_forward": (c_int, [c_int, c_int, c_int, c_void_p, c_void_p, c_float, c_void_p, c_void_p, c_void_p, c_void_p]),
    "htgs_l1_ssim_backward": (c_int, [c_int, c_int, c_int, c_void_p, c_void_p, c_void_p, c_float, c_void_p, c_void_p, c_void_p]),
    "htgs_dist2_workspace_bytes": (c_size_t, [c_int]),
    "htgs_dist2": (c_int, [c_int, c_void_p, c_void_p, c_void_p, c_size_t, c_void_p]),
}

EXPORTS = tuple(_SIGNATURES)
ABI_VERSION = 1
SORT_STATIC_MAX_PAIRS = 1 << 20  # HTGS_SORT_STATIC_MAX_PAIRS

GEOM_FIELDS = ("means2D", "conic_opacity", "feat", "clamped", "tiles_touched", "point_offsets", "block_sums", "num_rendered")
BIN_FIELDS = ("keys_sorted", "values_sorted", "keys_alt", "values_alt", "sort_scratch")
IMG_FIELDS = ("ranges", "n_contrib")

_lib = None


class HtgsError(RuntimeError):
    pass


def lib():
    """Load libhtgs.so once; raise loudly when it is absent (no CPU fallback)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise HtgsError(
                f"{LIB_PATH} not found: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                "(nvcc, sm_100a). This plugin has no CPU or PyTorch fallback.")
        l = ctypes.CDLL(LIB_PATH)
        for name, (res, args) in _SIGNATURES.items():
            fn = getattr(l, name)
            fn.restype = res
            fn.argtypes = args
        if l.htgs_abi_version() != ABI_VERSION:
            raise HtgsError(f"libhtgs.so ABI {l.htgs_abi_version()} != binding {ABI_VERSION}; rebuild")
        _lib = l
    return _lib


def check(status, what):
    if status != 0:
        raise HtgsError(f"{what}: {lib().htgs_last_error().decode(errors='replace')}")


def ptr(t):
    """Raw pointer of a tensor, NULL for None / empty placeholders (the reference passes empty
    tensors whose data_ptr() is 0, diff_gaussian_rasterization/__init__.py:200-210)."""
    if t is None or t.numel() == 0:
        return None
    return t.data_ptr()


def _layout(fn, names, *args):
    arr = (c_size_t * len(names))()
    n = fn(*args, arr, len(names))
    return {names[i]: int(arr[i]) for i in range(n)}


STAGES = ("preprocess", "scan", "make_keys", "sort", "tile_ranges", "blend_fwd", "zero_acc", "blend_bwd", "geom_bwd")


def profile_enable(on=True):
    lib().htgs_profile_enable(1 if on else 0)


def profile_read():
    """{stage: (total_ms, launches)} since the last read; waits for the recorded events."""
    ms = (ctypes.c_double * len(STAGES))()
    cnt = (c_int * len(STAGES))()
    if lib().htgs_profile_read(ms, cnt, len(STAGES)) < 0:
        raise HtgsError(lib().htgs_last_error().decode(errors="replace"))
    return {STAGES[i]: (float(ms[i]), int(cnt[i])) for i in range(len(STAGES))}


def geom_layout(P):
    return _layout(lib().htgs_geom_layout, GEOM_FIELDS, P)


def binning_layout(R, W, H):
    return _layout(lib().htgs_binning_layout, BIN_FIELDS, R, W, H)


def binning_capacity_from_bytes(nbytes, W, H):
    """Largest instance count whose binning buffer has exactly this size (buffers sized from a
    predicted count are laid out by that capacity, not by the true R)."""
    l = lib()
    lo, hi = 0, 1 << 30
    while lo < hi:  # largest c with bytes(c) <= nbytes
        mid = (lo + hi + 1) >> 1
        if l.htgs_binning_bytes(mid, W, H) <= nbytes:
            lo = mid
        else:
            hi = mid - 1
    return lo


def image_layout(W, H):
    return _layout(lib().htgs_image_layout, IMG_FIELDS, W, H)
