"""Opt-in rasterizer call on the model's STORED parameters.

The reference renderer builds its rasterizer inputs from the stored parameters on every call
(scene/gaussian_model_ht.py:127-188 getters, call site :834-866):

    scales    = torch.exp(pc._scaling)               rotations = F.normalize(pc._rotation)
    opacity   = torch.sigmoid(pc._opacity)           shs       = torch.cat((pc._features_dc, pc._features_rest), dim=1)

`rasterize_raw` takes `_xyz, _opacity, _features_dc, _features_rest, _scaling, _rotation` directly, applies
those activations inside the preprocess kernel and returns the same 4-tuple `(color, radii, depth, alpha)`;
its backward hands autograd the gradients with respect to the stored parameters.  The four intermediate
tensors, the 192 B/Gaussian concatenation and the split of its gradient are never materialised.  Replace

    out = rasterizer(means3D=pc.get_xyz, means2D=screenspace_points, shs=pc.get_features,
                     opacities=pc.get_opacity, scales=pc.get_scaling, rotations=pc.get_rotation)
by
    out = htgs_raw.rasterize_raw(pc.get_xyz, screenspace_points, pc._opacity, pc._features_dc,
                                 pc._features_rest, pc._scaling, pc._rotation, raster_settings)

Only the SH + scale/rotation configuration (the one HTGaussianRenderer trains with) has a raw form;
colors_precomp / cov3D_precomp callers keep using GaussianRasterizer.  CUDA only; no fallback.
"""
import ctypes

import torch

import htgs_native as nat
from diff_gaussian_rasterization import _C


def _check(t, name, shape_tail):
    if not t.is_cuda:
        raise RuntimeError(f"{name} must be a CUDA tensor: this rasterizer has no CPU path")
    if t.dtype != torch.float32:
        raise RuntimeError(f"expected scalar type Float but found {t.dtype} for {name}")
    if tuple(t.shape[1:]) != shape_tail:
        raise RuntimeError(f"{name} must have shape (P, {', '.join(map(str, shape_tail))}), got {tuple(t.shape)}")
    return t.contiguous()


def _grad_buffer(name, shape, align16, **opts):
    t = _C._arena_or_empty(name, shape, **opts)
    if align16 and t.data_ptr() % 16:
        t = torch.empty(shape, **opts)
    return t


class _RasterizeRaw(torch.autograd.Function):
    @staticmethod
    def forward(ctx, xyz, means2D, opacity_raw, features_dc, features_rest, scaling_raw, rotation_raw, raster_settings):
        rs = raster_settings
        if xyz.ndimension() != 2 or xyz.size(1) != 3:
            raise RuntimeError("means3D must have dimensions (num_points, 3)")
        xyz = _check(xyz, "xyz", (3,))
        P, H, W = int(xyz.size(0)), int(rs.image_height), int(rs.image_width)
        opacity_raw = _check(opacity_raw, "opacity", (1,))
        features_dc = _check(features_dc, "features_dc", (1, 3))
        features_rest = _check(features_rest, "features_rest", (15, 3))
        scaling_raw = _check(scaling_raw, "scaling", (3,))
        rotation_raw = _check(rotation_raw, "rotation", (4,))
        for t, n in ((opacity_raw, "opacity"), (features_dc, "features_dc"), (features_rest, "features_rest"),
                     (scaling_raw, "scaling"), (rotation_raw, "rotation")):
            if t.size(0) != P:
                raise RuntimeError(f"{n} has {t.size(0)} rows, xyz has {P}")
        dev = xyz.device
        f32 = dict(dtype=torch.float32, device=dev)
        u8 = dict(dtype=torch.uint8, device=dev)
        if P == 0:  # rasterize_points.cu:83
            ctx.empty = True
            return (torch.zeros((3, H, W), **f32), torch.zeros((0,), dtype=torch.int32, device=dev),
                    torch.zeros((1, H, W), **f32), torch.zeros((1, H, W), **f32))
        ctx.empty = False
        if features_rest.data_ptr() % 16:
            features_rest = features_rest.clone()  # fresh allocations are 512-byte aligned
        bg, view, proj, campos = (t.contiguous() for t in (rs.bg, rs.viewmatrix, rs.projmatrix, rs.campos))
        lib = nat.lib()
        with torch.cuda.device(dev):
            st = _C._stream(dev)
            color = torch.empty((3, H, W), **f32)
            depth = torch.empty((1, H, W), **f32)
            alpha = torch.empty((1, H, W), **f32)
            radii = torch.empty((P,), dtype=torch.int32, device=dev)
            geom = torch.empty(lib.htgs_geom_bytes(P), **u8)
            img = torch.empty(lib.htgs_image_bytes(W, H), **u8)
            pinned = _C._pinned_counter(dev)
            nat.check(lib.htgs_forward_preprocess_raw_async(
                P, int(rs.sh_degree), xyz.data_ptr(), scaling_raw.data_ptr(), float(rs.scale_modifier), rotation_raw.data_ptr(),
                opacity_raw.data_ptr(), features_dc.data_ptr(), features_rest.data_ptr(), view.data_ptr(), proj.data_ptr(),
                campos.data_ptr(), W, H, float(rs.tanfovx), float(rs.tanfovy), int(bool(rs.prefiltered)), geom.data_ptr(),
                radii.data_ptr(), pinned.data_ptr(), st), "htgs_forward_preprocess_raw_async")
            counted = torch.cuda.Event()
            counted.record(torch.cuda.current_stream(dev))
            out_args = (img.data_ptr(), color.data_ptr(), depth.data_ptr(), alpha.data_ptr(), st)
            key = (dev.index, P, W, H)
            predicted = _C._R_PREDICTION.get(key) if _C.SPECULATIVE and not rs.debug else None
            R = None
            carve = None  # the instance count the binning buffer ends up carved with (see _C._Rendered)
            if predicted is not None:
                capacity = int(predicted * _C.CAPACITY_MARGIN) + 8192
                binning = torch.empty(lib.htgs_binning_bytes(capacity, W, H), **u8)
                nat.check(lib.htgs_forward_render_capacity(P, capacity, W, H, bg.data_ptr(), radii.data_ptr(), geom.data_ptr(),
                                                           binning.data_ptr(), *out_args), "htgs_forward_render_capacity")
                counted.synchronize()
                R = int(pinned[0]) & 0xFFFFFFFF
                carve = capacity
            if predicted is None or R > capacity:
                counted.synchronize()
                R = int(pinned[0]) & 0xFFFFFFFF
                carve = R
                binning = torch.empty(lib.htgs_binning_bytes(R, W, H) if R > 0 else 0, **u8)
                nat.check(lib.htgs_forward_render(P, R, W, H, bg.data_ptr(), radii.data_ptr(), geom.data_ptr(),
                                                  nat.ptr(binning), *out_args), "htgs_forward_render")
            _C._R_PREDICTION[key] = R
            if rs.debug:
                torch.cuda.synchronize(dev)
        ctx.raster_settings = rs
        ctx.num_rendered = _C._Rendered(R, carve)
        ctx.save_for_backward(xyz, opacity_raw, features_dc, features_rest, scaling_raw, rotation_raw, radii, geom, binning,
                              img, alpha)
        ctx.mark_non_differentiable(radii)
        return color, radii, depth, alpha

    @staticmethod
    def backward(ctx, grad_color, grad_radii, grad_depth, grad_alpha):
        if ctx.empty:
            return (None,) * 8
        rs = ctx.raster_settings
        xyz, opacity_raw, features_dc, features_rest, scaling_raw, rotation_raw, radii, geom, binning, img, alpha = ctx.saved_tensors
        dev = xyz.device
        P, H, W = int(xyz.size(0)), int(rs.image_height), int(rs.image_width)
        f32 = dict(dtype=torch.float32, device=dev)
        grad_color, grad_depth, grad_alpha = (g.contiguous() for g in (grad_color, grad_depth, grad_alpha))
        bg, view, proj, campos = (t.contiguous() for t in (rs.bg, rs.viewmatrix, rs.projmatrix, rs.campos))
        lib = nat.lib()
        with torch.cuda.device(dev):
            st = _C._stream(dev)
            dmeans2D = torch.empty((P, 3), **f32)
            dxyz = _grad_buffer("means3D", (P, 3), False, **f32)
            dopacity = _grad_buffer("opacities", (P, 1), False, **f32)
            dfdc = _grad_buffer("features_dc", (P, 1, 3), False, **f32)
            dfrest = _grad_buffer("features_rest", (P, 15, 3), True, **f32)
            dscaling = _grad_buffer("scales", (P, 3), False, **f32)
            drot = _grad_buffer("rotations", (P, 4), True, **f32)
            scratch = torch.empty(lib.htgs_backward_scratch_bytes(P), dtype=torch.uint8, device=dev)
            nat.check(lib.htgs_backward_raw(
                P, int(rs.sh_degree), int(ctx.num_rendered.carve), W, H, bg.data_ptr(), xyz.data_ptr(), features_dc.data_ptr(),
                features_rest.data_ptr(), alpha.data_ptr(), scaling_raw.data_ptr(), float(rs.scale_modifier),
                rotation_raw.data_ptr(), opacity_raw.data_ptr(), view.data_ptr(), proj.data_ptr(), campos.data_ptr(),
                float(rs.tanfovx), float(rs.tanfovy), radii.data_ptr(), geom.data_ptr(), nat.ptr(binning), img.data_ptr(),
                grad_color.data_ptr(), grad_depth.data_ptr(), grad_alpha.data_ptr(), scratch.data_ptr(), dmeans2D.data_ptr(),
                dxyz.data_ptr(), dfdc.data_ptr(), dfrest.data_ptr(), dopacity.data_ptr(), dscaling.data_ptr(), drot.data_ptr(),
                st), "htgs_backward_raw")
            if rs.debug:
                torch.cuda.synchronize(dev)
        return dxyz, dmeans2D, dopacity, dfdc, dfrest, dscaling, drot, None


def rasterize_raw(xyz, means2D, opacity_raw, features_dc, features_rest, scaling_raw, rotation_raw, raster_settings):
    """(color[3,H,W], radii int32[P], depth[1,H,W], alpha[1,H,W]) from the stored parameters; see module docstring."""
    return _RasterizeRaw.apply(xyz, means2D, opacity_raw, features_dc, features_rest, scaling_raw, rotation_raw, raster_settings)
