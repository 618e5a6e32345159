"""Drop-in `diff_gaussian_rasterization` plugin backed by the B200-native library.

Mirrors the public surface of the reference plugin
(submodules/diff-gaussian-rasterization/diff_gaussian_rasterization/__init__.py:
cpu_deep_copy_tuple :17, rasterize_gaussians :21, _RasterizeGaussians :44,
GaussianRasterizationSettings :160, GaussianRasterizer :174) so that
gaussian_renderer/__init__.py:38-96 and scene/gaussian_model_ht.py:809-880 run
unchanged: same names, same argument order, same 4-tuple
(color, radii, depth, alpha), same gradient slots.
"""
from typing import NamedTuple

import torch
import torch.nn as nn

from . import _C


class GaussianRasterizationSettings(NamedTuple):
    """Per-call camera/config tuple (reference __init__.py:160-172). Matrices are in the
    row-vector convention of scene/cameras.py:78-97 (stored transposed)."""
    image_height: int
    image_width: int
    tanfovx: float
    tanfovy: float
    bg: torch.Tensor
    scale_modifier: float
    viewmatrix: torch.Tensor
    projmatrix: torch.Tensor
    sh_degree: int
    campos: torch.Tensor
    prefiltered: bool
    debug: bool


def cpu_deep_copy_tuple(input_tuple):
    """Host-side snapshot of an argument tuple, taken before a debug-mode call can corrupt it."""
    return tuple(x.cpu().clone() if isinstance(x, torch.Tensor) else x for x in input_tuple)


def _call_with_snapshot(fn, args, debug, dump_name, phase):
    """debug=True keeps a CPU copy of the inputs and writes it to `dump_name` if the native call
    throws (reference __init__.py:83-90, :135-142)."""
    if not debug:
        return fn(*args)
    saved = cpu_deep_copy_tuple(args)
    try:
        return fn(*args)
    except Exception:
        torch.save(saved, dump_name)
        print(f"\nAn error occured in {phase}. Please forward {dump_name} for debugging.")
        raise


class _RasterizeGaussians(torch.autograd.Function):
    @staticmethod
    def forward(ctx, means3D, means2D, sh, colors_precomp, opacities, scales, rotations, cov3Ds_precomp,
                raster_settings):
        rs = raster_settings
        native_args = (rs.bg, means3D, colors_precomp, opacities, scales, rotations, rs.scale_modifier,
                       cov3Ds_precomp, rs.viewmatrix, rs.projmatrix, rs.tanfovx, rs.tanfovy, rs.image_height,
                       rs.image_width, sh, rs.sh_degree, rs.campos, rs.prefiltered, rs.debug)
        (num_rendered, color, depth, alpha, radii, geom_buffer, binning_buffer,
         img_buffer) = _call_with_snapshot(_C.rasterize_gaussians, native_args, rs.debug, "snapshot_fw.dump", "forward")
        ctx.raster_settings = rs
        ctx.num_rendered = num_rendered
        # color and depth are NOT saved: callers modify them in place
        # (trainer/ht3dgs_trainer.py:1292, trainer/losses.py:116-117).
        ctx.save_for_backward(colors_precomp, means3D, scales, rotations, cov3Ds_precomp, radii, sh, geom_buffer,
                              binning_buffer, img_buffer, alpha)
        return color, radii, depth, alpha

    @staticmethod
    def backward(ctx, grad_color, grad_radii, grad_depth, grad_alpha):
        rs = ctx.raster_settings
        (colors_precomp, means3D, scales, rotations, cov3Ds_precomp, radii, sh, geom_buffer, binning_buffer,
         img_buffer, alpha) = ctx.saved_tensors
        native_args = (rs.bg, means3D, radii, colors_precomp, scales, rotations, rs.scale_modifier, cov3Ds_precomp,
                       rs.viewmatrix, rs.projmatrix, rs.tanfovx, rs.tanfovy, grad_color, grad_depth, grad_alpha, sh,
                       rs.sh_degree, rs.campos, geom_buffer, ctx.num_rendered, binning_buffer, img_buffer, alpha,
                       rs.debug)
        (grad_means2D, grad_colors_precomp, grad_opacities, grad_means3D, grad_cov3Ds_precomp, grad_sh, grad_scales,
         grad_rotations) = _call_with_snapshot(_native_backward, native_args, rs.debug, "snapshot_bw.dump", "backward")
        # slot order = forward's input order (reference __init__.py:146-156)
        return (grad_means3D, grad_means2D, grad_sh, grad_colors_precomp, grad_opacities, grad_scales, grad_rotations,
                grad_cov3Ds_precomp, None)


def _native_backward(*args):
    # Gradients of the empty placeholder inputs cannot be consumed by autograd; skip producing them.
    return _C.backward_impl(*args, mirror_all=False)


def rasterize_gaussians(means3D, means2D, sh, colors_precomp, opacities, scales, rotations, cov3Ds_precomp,
                        raster_settings):
    return _RasterizeGaussians.apply(means3D, means2D, sh, colors_precomp, opacities, scales, rotations,
                                     cov3Ds_precomp, raster_settings)


class GaussianRasterizer(nn.Module):
    def __init__(self, raster_settings):
        super().__init__()
        self.raster_settings = raster_settings

    def markVisible(self, positions):
        """bool[P]: which points pass the near-plane test of the current camera."""
        with torch.no_grad():
            rs = self.raster_settings
            return _C.mark_visible(positions, rs.viewmatrix, rs.projmatrix)

    def forward(self, means3D, means2D, opacities, shs=None, colors_precomp=None, scales=None, rotations=None,
                cov3D_precomp=None):
        if (shs is None) == (colors_precomp is None):
            raise Exception('Please provide excatly one of either SHs or precomputed colors!')
        have_sr = scales is not None or rotations is not None
        if ((scales is None or rotations is None) and cov3D_precomp is None) or (have_sr and cov3D_precomp is not None):
            raise Exception('Please provide exactly one of either scale/rotation pair or precomputed 3D covariance!')

        # the reference hands empty CPU tensors to the native side for absent options (:200-210)
        def or_empty(t):
            return torch.Tensor([]) if t is None else t

        return rasterize_gaussians(means3D, means2D, or_empty(shs), or_empty(colors_precomp), opacities,
                                   or_empty(scales), or_empty(rotations), or_empty(cov3D_precomp),
                                   self.raster_settings)
