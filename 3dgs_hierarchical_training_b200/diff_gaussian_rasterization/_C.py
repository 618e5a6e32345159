"""Stand-in for the reference's pybind11 module `diff_gaussian_rasterization._C`
(submodules/diff-gaussian-rasterization/ext.cpp:15-19, rasterize_points.cu).

Same three entry points, same argument order, same returned tuples -- but the
work is done by libhtgs.so through the C ABI of include/htgs.h; this file only
allocates torch tensors and marshals raw pointers.  No fallback path exists:
CPU tensors or a missing library raise.
"""
import ctypes

import torch

import htgs_native as nat

NUM_CHANNELS = 3  # cuda_rasterizer/config.h:15

# Speculative sizing of the binning buffer (htgs_forward_render_capacity): the previous call's R for the
# same (device, P, W, H) predicts the next one, so the whole forward is enqueued without the host
# waiting for R; the exact R is read afterwards while the GPU is already blending.
#
# This is the one piece of process-global state of the plugin mirror: `_R_PREDICTION` (a dict keyed by
# (device, P, W, H)) and ONE pinned 4-byte counter per device (`_PINNED_R`) that the async preprocess writes R into.
# A forward reads its own R back from that counter before it returns, so forwards issued one after another from one
# host thread -- the reference trainer's pattern -- are safe on any stream.  Two forwards IN FLIGHT at the same time on
# one device (two host threads, or two streams driven without waiting for the first call to return) would share the
# counter and may read each other's R: give each thread its own process, or set SPECULATIVE = False (the reference's
# blocking two-phase flow, no shared state).
_R_PREDICTION = {}
_PINNED_R = {}
SPECULATIVE = True
CAPACITY_MARGIN = 1.25


# Optional gradient arena (view-parallel training): when set, the backward writes the five parameter
# gradients straight into slices of one flat buffer (the one that is all-reduced) instead of fresh tensors,
# which removes the 236 B/Gaussian pack copy before the collective.  Only meaningful when every parameter
# receives one backward per step; see htgs_dist.FlatGradReducer.
_GRAD_ARENA = None


def set_grad_arena(flat, slices, owners=None):
    """flat: 1-D float32 CUDA tensor; slices: {"means3D"|"shs"|"opacities"|"scales"|"rotations": (lo, hi)}.
    owners (optional): {name: leaf tensor whose .grad adopts the slice}.  When a leaf still holds last step's
    arena slice as its .grad (zero_grad(set_to_none=False), or a second view rendered before the optimizer step) the
    backward must not overwrite that memory: autograd would then add the new gradient to itself.  With owners given,
    such a backward gets ordinary fresh tensors and autograd accumulates them INTO the arena slice, which is the
    correct sum.  Without owners the caller promises one backward per parameter per step and .grad = None in between.
    Pass (None, None) to clear."""
    global _GRAD_ARENA
    _GRAD_ARENA = None if flat is None else (flat, dict(slices), dict(owners) if owners else None)


def _arena_or_empty(name, shape, **opts):
    if _GRAD_ARENA is not None:
        flat, slices, owners = _GRAD_ARENA
        rng = slices.get(name)
        n = 1
        for v in shape:
            n *= v
        if rng is not None and rng[1] - rng[0] == n and flat.device == opts["device"]:
            owner = owners.get(name) if owners else None
            held = owner is not None and owner.grad is not None and owner.grad.data_ptr() == flat.data_ptr() + 4 * rng[0]
            # dL_drotations is written with 128-bit stores: a slice that does not start 16-byte aligned (P not a multiple
            # of 4 in the flat 59-float layout) gets an ordinary tensor and is packed by the reducer afterwards
            misaligned = name == "rotations" and (flat.data_ptr() + 4 * rng[0]) % 16 != 0
            if not held and not misaligned:
                return flat[rng[0]:rng[1]].view(shape)  # a fresh view object: autograd may adopt it as .grad
    return torch.empty(shape, **opts)


# Optional compact SH mode (view-parallel training, htgs_dist.CompactSHGradReducer): when set, the backward of the SH
# path does not produce dL_dsh at all.  It writes the clamp-masked colour gradient and the camera position of the view
# into the slot the provider names (htgs_backward_compact) and returns None for the SH gradient; the reducer rebuilds
# the summed dL_dsh from the slots of all views on every rank.
_SH_COMPACT = None


def set_sh_compact(provider):
    """provider(device) -> (slot_ptr: int, is_multicast: bool) for the next backward on that device; None clears."""
    global _SH_COMPACT
    _SH_COMPACT = provider


class _Rendered(int):
    """`num_rendered` as the reference returns it (an int), carrying the instance count the binning buffer of that
    forward was CARVED with (the exact R, or the predicted capacity of the speculative path).  The backward hands that
    count to the C ABI so that both sides see the same layout of the binning buffer -- which is what lets it use the
    reach masks the forward blend cached there instead of evaluating them again."""

    def __new__(cls, value, carve):
        obj = super().__new__(cls, value)
        obj.carve = int(carve)
        return obj


def _pinned_counter(dev):
    t = _PINNED_R.get(dev.index)
    if t is None:
        t = torch.zeros(1, dtype=torch.int32).pin_memory()
        _PINNED_R[dev.index] = t
    return t


def _f32c(t):
    """What `.contiguous().data<float>()` does in rasterize_points.cu:96-115."""
    if t.numel() == 0:
        return t
    if t.dtype != torch.float32:
        raise RuntimeError(f"expected scalar type Float but found {t.dtype}")
    return t.contiguous()


def _require_cuda(t, name):
    if not t.is_cuda:
        raise RuntimeError(f"{name} must be a CUDA tensor: this rasterizer has no CPU path")


def _stream(device):
    return ctypes.c_void_p(torch.cuda.current_stream(device).cuda_stream)


def rasterize_gaussians(background, means3D, colors, opacity, scales, rotations, scale_modifier, cov3D_precomp,
                        viewmatrix, projmatrix, tan_fovx, tan_fovy, image_height, image_width, sh, degree, campos,
                        prefiltered, debug):
    """RasterizeGaussiansCUDA (rasterize_points.cu:35-119).

    Returns (num_rendered, out_color[3,H,W], out_depth[1,H,W], out_alpha[1,H,W], radii[P] int32,
    geomBuffer, binningBuffer, imgBuffer) with the three buffers as opaque uint8 tensors."""
    if means3D.ndimension() != 2 or means3D.size(1) != 3:
        raise RuntimeError("means3D must have dimensions (num_points, 3)")
    _require_cuda(means3D, "means3D")
    dev = means3D.device
    P, H, W = int(means3D.size(0)), int(image_height), int(image_width)
    f32 = dict(dtype=torch.float32, device=dev)
    u8 = dict(dtype=torch.uint8, device=dev)
    if P == 0:
        # rasterize_points.cu:83: nothing is launched; images stay zero (background is NOT applied)
        return (0, torch.zeros((NUM_CHANNELS, H, W), **f32), torch.zeros((1, H, W), **f32), torch.zeros((1, H, W), **f32),
                torch.zeros((0,), dtype=torch.int32, device=dev), torch.empty(0, **u8), torch.empty(0, **u8),
                torch.empty(0, **u8))
    M = int(sh.size(1)) if sh.numel() != 0 and sh.size(0) != 0 else 0
    means3D, colors, opacity, scales, rotations, cov3D_precomp, sh = map(
        _f32c, (means3D, colors, opacity, scales, rotations, cov3D_precomp, sh))
    background, viewmatrix, projmatrix, campos = map(_f32c, (background, viewmatrix, projmatrix, campos))
    lib = nat.lib()
    with torch.cuda.device(dev):
        st = _stream(dev)
        out_color = torch.empty((NUM_CHANNELS, H, W), **f32)
        out_depth = torch.empty((1, H, W), **f32)
        out_alpha = torch.empty((1, H, W), **f32)
        radii = torch.empty((P,), dtype=torch.int32, device=dev)
        geom = torch.empty(lib.htgs_geom_bytes(P), **u8)
        img = torch.empty(lib.htgs_image_bytes(W, H), **u8)
        pre_args = (P, int(degree), M, nat.ptr(means3D), nat.ptr(scales), float(scale_modifier), nat.ptr(rotations),
                    nat.ptr(opacity), nat.ptr(sh), nat.ptr(cov3D_precomp), nat.ptr(colors), nat.ptr(viewmatrix),
                    nat.ptr(projmatrix), nat.ptr(campos), W, H, float(tan_fovx), float(tan_fovy), int(bool(prefiltered)),
                    geom.data_ptr(), radii.data_ptr())
        out_args = (img.data_ptr(), out_color.data_ptr(), out_depth.data_ptr(), out_alpha.data_ptr(), st)
        key = (dev.index, P, W, H)
        predicted = _R_PREDICTION.get(key) if SPECULATIVE and not debug else None
        if predicted is None:
            # first call for this shape: the reference's two-phase flow, host waits for R once
            rendered = ctypes.c_int64(0)
            nat.check(lib.htgs_forward_preprocess(*pre_args, ctypes.byref(rendered), st), "htgs_forward_preprocess")
            R = int(rendered.value)
            carve = R
            binning = torch.empty(lib.htgs_binning_bytes(R, W, H) if R > 0 else 0, **u8)
            nat.check(lib.htgs_forward_render(P, R, W, H, nat.ptr(background), radii.data_ptr(), geom.data_ptr(),
                                              nat.ptr(binning), *out_args), "htgs_forward_render")
        else:
            capacity = int(predicted * CAPACITY_MARGIN) + 8192
            pinned = _pinned_counter(dev)
            nat.check(lib.htgs_forward_preprocess_async(*pre_args, pinned.data_ptr(), st), "htgs_forward_preprocess_async")
            counted = torch.cuda.Event()
            counted.record(torch.cuda.current_stream(dev))
            binning = torch.empty(lib.htgs_binning_bytes(capacity, W, H), **u8)
            nat.check(lib.htgs_forward_render_capacity(P, capacity, W, H, nat.ptr(background), radii.data_ptr(),
                                                       geom.data_ptr(), binning.data_ptr(), *out_args),
                      "htgs_forward_render_capacity")
            counted.synchronize()  # the GPU is already sorting/blending while the host reads R
            R = int(pinned[0]) & 0xFFFFFFFF
            carve = capacity
            if R > capacity:  # prediction too small: redo phase 2 with the exact size
                carve = R
                binning = torch.empty(lib.htgs_binning_bytes(R, W, H), **u8)
                nat.check(lib.htgs_forward_render(P, R, W, H, nat.ptr(background), radii.data_ptr(), geom.data_ptr(),
                                                  binning.data_ptr(), *out_args), "htgs_forward_render")
        _R_PREDICTION[key] = R
        if debug:
            torch.cuda.synchronize(dev)  # CHECK_CUDA(..., debug) semantics, auxiliary.h:166-173
    return _Rendered(R, carve), out_color, out_depth, out_alpha, radii, geom, binning, img


def backward_impl(background, means3D, radii, colors, scales, rotations, scale_modifier, cov3D_precomp, viewmatrix,
                  projmatrix, tan_fovx, tan_fovy, dL_dout_color, dL_dout_depth, dL_dout_alpha, sh, degree, campos,
                  geomBuffer, R, binningBuffer, imageBuffer, alphas, debug, mirror_all=True):
    """Core of the backward.  With mirror_all=False the gradients nobody can consume (dL_dcolors on the
    SH path, dL_dcov3D on the scale/rotation path) are neither computed into memory nor allocated and
    come back as None."""
    _require_cuda(means3D, "means3D")
    dev = means3D.device
    P = int(means3D.size(0))
    H, W = int(dL_dout_color.size(1)), int(dL_dout_color.size(2))
    M = int(sh.size(1)) if sh.numel() != 0 and sh.size(0) != 0 else 0
    # the count the forward carved the binning buffer with (a plain int from another caller means the exact R)
    R = getattr(R, "carve", int(R))
    f32 = dict(dtype=torch.float32, device=dev)
    has_sh, has_colors = sh.numel() != 0, colors.numel() != 0
    has_scales, has_cov = scales.numel() != 0, cov3D_precomp.numel() != 0
    if P == 0:
        z = lambda *s: torch.zeros(s, **f32)
        return z(0, 3), z(0, NUM_CHANNELS), z(0, 1), z(0, 3), z(0, 6), z(0, M, 3), z(0, 3), z(0, 4)
    means3D, colors, scales, rotations, cov3D_precomp, sh, alphas = map(
        _f32c, (means3D, colors, scales, rotations, cov3D_precomp, sh, alphas))
    background, viewmatrix, projmatrix, campos = map(_f32c, (background, viewmatrix, projmatrix, campos))
    dL_dout_color, dL_dout_depth, dL_dout_alpha = map(_f32c, (dL_dout_color, dL_dout_depth, dL_dout_alpha))
    radii = radii.contiguous()
    lib = nat.lib()
    with torch.cuda.device(dev):
        st = _stream(dev)
        dmeans2D = torch.empty((P, 3), **f32)
        dopacity = _arena_or_empty("opacities", (P, 1), **f32)
        dmeans3D = _arena_or_empty("means3D", (P, 3), **f32)
        want_colors = has_colors or mirror_all
        want_cov = has_cov or mirror_all
        dcolors = torch.empty((P, NUM_CHANNELS), **f32) if want_colors else None
        dcov3D = torch.empty((P, 6), **f32) if want_cov else None
        compact = _SH_COMPACT is not None and has_sh and not has_colors
        if compact:
            dsh = None
        else:
            dsh = _arena_or_empty("shs", (P, M, 3), **f32) if has_sh else torch.zeros((P, M, 3), **f32)
        dscales = _arena_or_empty("scales", (P, 3), **f32) if has_scales else torch.zeros((P, 3), **f32)
        drot = _arena_or_empty("rotations", (P, 4), **f32) if has_scales else torch.zeros((P, 4), **f32)
        scratch = torch.empty(lib.htgs_backward_scratch_bytes(P), dtype=torch.uint8, device=dev)
        if compact:
            slot_ptr, slot_mc = _SH_COMPACT(dev)
            nat.check(lib.htgs_backward_compact(
                P, int(degree), M, int(R), W, H, nat.ptr(background), nat.ptr(means3D), nat.ptr(sh), nat.ptr(alphas),
                nat.ptr(scales), float(scale_modifier), nat.ptr(rotations), nat.ptr(cov3D_precomp), nat.ptr(viewmatrix),
                nat.ptr(projmatrix), nat.ptr(campos), float(tan_fovx), float(tan_fovy), radii.data_ptr(), geomBuffer.data_ptr(),
                nat.ptr(binningBuffer), imageBuffer.data_ptr(), nat.ptr(dL_dout_color), nat.ptr(dL_dout_depth),
                nat.ptr(dL_dout_alpha), scratch.data_ptr(), dmeans2D.data_ptr(), dopacity.data_ptr(), dmeans3D.data_ptr(),
                nat.ptr(dcov3D), nat.ptr(dscales) if has_scales else None, nat.ptr(drot) if has_scales else None,
                ctypes.c_void_p(int(slot_ptr)), int(bool(slot_mc)), st), "htgs_backward_compact")
            if debug:
                torch.cuda.synchronize(dev)
            return dmeans2D, dcolors, dopacity, dmeans3D, dcov3D, dsh, dscales, drot
        nat.check(lib.htgs_backward(
            P, int(degree), M, int(R), W, H, nat.ptr(background), nat.ptr(means3D), nat.ptr(sh), nat.ptr(colors),
            nat.ptr(alphas), nat.ptr(scales), float(scale_modifier), nat.ptr(rotations), nat.ptr(cov3D_precomp),
            nat.ptr(viewmatrix), nat.ptr(projmatrix), nat.ptr(campos), float(tan_fovx), float(tan_fovy),
            radii.data_ptr(), geomBuffer.data_ptr(), nat.ptr(binningBuffer), imageBuffer.data_ptr(),
            nat.ptr(dL_dout_color), nat.ptr(dL_dout_depth), nat.ptr(dL_dout_alpha), scratch.data_ptr(),
            dmeans2D.data_ptr(), dopacity.data_ptr(), nat.ptr(dcolors), dmeans3D.data_ptr(), nat.ptr(dcov3D),
            nat.ptr(dsh) if has_sh else None, nat.ptr(dscales) if has_scales else None,
            nat.ptr(drot) if has_scales else None, st), "htgs_backward")
        if debug:
            torch.cuda.synchronize(dev)
    return dmeans2D, dcolors, dopacity, dmeans3D, dcov3D, dsh, dscales, drot


def rasterize_gaussians_backward(background, means3D, radii, colors, scales, rotations, scale_modifier, cov3D_precomp,
                                 viewmatrix, projmatrix, tan_fovx, tan_fovy, dL_dout_color, dL_dout_depth,
                                 dL_dout_alpha, sh, degree, campos, geomBuffer, R, binningBuffer, imageBuffer, alphas,
                                 debug):
    """RasterizeGaussiansBackwardCUDA (rasterize_points.cu:121-208).

    Returns (dL_dmeans2D[P,3], dL_dcolors[P,3], dL_dopacity[P,1], dL_dmeans3D[P,3], dL_dcov3D[P,6],
    dL_dsh[P,M,3], dL_dscales[P,3], dL_drotations[P,4])."""
    return backward_impl(background, means3D, radii, colors, scales, rotations, scale_modifier, cov3D_precomp,
                         viewmatrix, projmatrix, tan_fovx, tan_fovy, dL_dout_color, dL_dout_depth, dL_dout_alpha, sh,
                         degree, campos, geomBuffer, R, binningBuffer, imageBuffer, alphas, debug, mirror_all=True)


def mark_visible(means3D, viewmatrix, projmatrix):
    """markVisible (rasterize_points.cu:210-229): bool[P], view-space z > 0.2."""
    _require_cuda(means3D, "means3D")
    dev = means3D.device
    P = int(means3D.size(0))
    present = torch.zeros((P,), dtype=torch.bool, device=dev)
    if P != 0:
        with torch.cuda.device(dev):
            nat.check(nat.lib().htgs_mark_visible(P, nat.ptr(_f32c(means3D)), nat.ptr(_f32c(viewmatrix)),
                                                  nat.ptr(_f32c(projmatrix)), present.data_ptr(), _stream(dev)),
                      "htgs_mark_visible")
    return present
