"""Merge-time operations of the hierarchical trainer on this library's kernels (opt-in).

    calc_importance(gs_render, cameras, pipe)        trainer/ht3dgs_trainer.py:1428-1462
    importance_prune_mask(importance, prune_ratio)   the amax / topk(largest=False) / mask lines of merge_two_3DGS (:232-238)
    merge_two_models(dst, src, T, mask_dst, mask_src) the rest of merge_two_3DGS (:240-270): prune dst, transform + filter src, append
    capture(model) / restore(model, args, opt)        scene/gaussian_model_ht.py:90-124, with FusedAdam kept as the optimizer

The model objects are the reference's own (HTGaussianModel / CF3DGS_Render): attributes are read and written by the names
the reference uses, so the trainer can call these in place of its methods.  CUDA only; no fallback."""
import ctypes
import math

import torch

import htgs_native as nat
from htgs_optim import FusedAdam, _stream, prune_optimizer

GROUPS = ("xyz", "f_dc", "f_rest", "opacity", "scaling", "rotation")
_ATTR = {"xyz": "_xyz", "f_dc": "_features_dc", "f_rest": "_features_rest", "opacity": "_opacity", "scaling": "_scaling",
         "rotation": "_rotation"}


def accumulate_importance(importance, state, sh_degree, means3D, campos, background, dL_dcolor):
    """One view: importance[P, M, 3] += |dL/dsh| for the upstream colour gradient `dL_dcolor` ([3,H,W]; depth and alpha get
    none).  `state` = the tuple `_C.rasterize_gaussians` returned for that view (R, color, depth, alpha, radii, geom,
    binning, img)."""
    R, _, _, alpha, radii, geom, binning, img = state
    P, M = importance.shape[0], importance.shape[1]
    H, W = dL_dcolor.shape[1], dL_dcolor.shape[2]
    dev = importance.device
    lib = nat.lib()
    with torch.cuda.device(dev):
        scratch = torch.empty(lib.htgs_backward_scratch_bytes(P), dtype=torch.uint8, device=dev)
        nat.check(lib.htgs_importance_accumulate(P, int(sh_degree), M, int(getattr(R, "carve", R)), W, H, nat.ptr(background.contiguous()),
                                                 means3D.contiguous().data_ptr(), campos.contiguous().data_ptr(), alpha.data_ptr(),
                                                 radii.data_ptr(), geom.data_ptr(), nat.ptr(binning), img.data_ptr(),
                                                 dL_dcolor.contiguous().data_ptr(), scratch.data_ptr(), importance.data_ptr(),
                                                 _stream(dev)), "htgs_importance_accumulate")


@torch.no_grad()
def calc_importance(gs_render, cameras, pipe):
    """calc_importance of the reference: sum over the cameras of |d(image.sum())/d(features)| / number of pixels, [P, 48].

    The reference renders each camera through autograd with abs() hooks on the feature tensors.  Here each camera costs a
    forward, the blend backward and one streaming kernel: `image` is `clamp(colour, 0, 1)` (gaussian_model_ht.py:883), so
    the loss gradient is 1 where 0 <= colour <= 1 and 0 elsewhere, and dL/dsh is formed in place as basis (x) colour gradient."""
    from diff_gaussian_rasterization import _C
    if getattr(pipe, "convert_SHs_python", False) or getattr(pipe, "compute_cov3D_python", False):
        raise RuntimeError("calc_importance: only the native SH / scale+rotation path is implemented (pipe.convert_SHs_python "
                           "and pipe.compute_cov3D_python must be False, as train_leaf / train_nonleaf set them)")
    g = gs_render.gaussians
    means3D, opacity, scales, rotations, shs = g.get_xyz, g.get_opacity, g.get_scaling, g.get_rotation, g.get_features
    P, M = shs.shape[0], shs.shape[1]
    importance = torch.zeros((P, M, 3), dtype=torch.float32, device=means3D.device)
    empty = torch.Tensor([])
    num_pixels = 0
    for cam in cameras:
        H, W = int(cam.image_height), int(cam.image_width)
        state = _C.rasterize_gaussians(gs_render.bg_color, means3D, empty, opacity, scales, rotations, 1.0, empty,
                                       cam.world_view_transform, cam.full_proj_transform, math.tan(cam.FoVx * 0.5),
                                       math.tan(cam.FoVy * 0.5), H, W, shs, g.active_sh_degree, cam.camera_center, False, False)
        color = state[1]
        dL_dcolor = ((color >= 0.0) & (color <= 1.0)).to(torch.float32)
        accumulate_importance(importance, state, g.active_sh_degree, means3D, cam.camera_center, gs_render.bg_color, dL_dcolor)
        num_pixels += H * W
    return (importance / num_pixels).flatten(-2)


def importance_prune_mask(importance, prune_ratio):
    """bool[P]: True for the int(P * prune_ratio) Gaussians with the smallest importance.amax(-1) -- what
    `topk(importance.amax(-1), k, largest=False)` + the mask fill of merge_two_3DGS select.  Ties go to the lower index."""
    if importance.dim() != 2 or not importance.is_cuda or importance.dtype != torch.float32:
        raise RuntimeError("importance must be a float32 [P, n] CUDA tensor")
    P, n_per = importance.shape
    k = int(P * prune_ratio)
    dev = importance.device
    mask = torch.empty(P, dtype=torch.uint8, device=dev)
    if P == 0:
        return mask.bool()
    lib = nat.lib()
    with torch.cuda.device(dev):
        ws = torch.empty(lib.htgs_importance_select_workspace_bytes(P), dtype=torch.uint8, device=dev)
        nat.check(lib.htgs_importance_select(P, n_per, importance.contiguous().data_ptr(), k, ws.data_ptr(), mask.data_ptr(), _stream(dev)),
                  "htgs_importance_select")
    return mask.bool()


def transform_points(xyz, transform_matrix):
    """[x 1] @ T.T, divided by the homogeneous coordinate (merge_two_3DGS :254-257); T is a 4x4 tensor (source -> destination)."""
    dev = xyz.device
    P = xyz.shape[0]
    out = torch.empty((P, 3), dtype=torch.float32, device=dev)
    T = transform_matrix.detach().to(device=dev, dtype=torch.float32).contiguous()
    with torch.cuda.device(dev):
        nat.check(nat.lib().htgs_transform_points(P, xyz.detach().contiguous().data_ptr(), T.data_ptr(), out.data_ptr(), _stream(dev)),
                  "htgs_transform_points")
    return out


def _set_model_params(model, new_params):
    for name, attr in _ATTR.items():
        if name in new_params:
            setattr(model, attr, new_params[name])


def prune_model(model, mask):
    """HTGaussianModel.prune_points (gaussian_model_ht.py:573-588): one mask scan and one launch for the six parameters,
    their twelve Adam moments and the three statistics tensors."""
    keep = ~mask
    new_params, extra = prune_optimizer(model.optimizer, keep, extra=(model.xyz_gradient_accum, model.denom, model.max_radii2D))
    _set_model_params(model, new_params)
    model.xyz_gradient_accum, model.denom, model.max_radii2D = extra


def append_rows(model, src_tensors, keep_src):
    """densification_postfix (gaussian_model_ht.py:613-628) with the new rows taken from `src_tensors[name][keep_src]`:
    every parameter becomes cat(param, src[keep]), both Adam moments get as many zero rows, the densification statistics
    are reset to zeros of the new size.  The rows of all six source tensors are compacted by ONE mask scan and ONE launch,
    straight into the tail of the enlarged destination tensors (no intermediate src[keep] tensors)."""
    opt = model.optimizer
    dev = keep_src.device
    keep = keep_src.contiguous()
    if keep.dtype not in (torch.bool, torch.uint8) or keep.dim() != 1:
        raise RuntimeError("keep_src must be a 1-D bool/uint8 tensor")
    P_src = int(keep.numel())
    lib = nat.lib()
    with torch.cuda.device(dev):
        st = _stream(dev)
        scratch = torch.empty(lib.htgs_row_scan_scratch_bytes(P_src), dtype=torch.uint8, device=dev)
        pinned = torch.zeros(1, dtype=torch.int32).pin_memory()
        nat.check(lib.htgs_row_scan(P_src, keep.data_ptr(), scratch.data_ptr(), pinned.data_ptr(), st), "htgs_row_scan")
        torch.cuda.current_stream(dev).synchronize()  # the one host read: the new size depends on it
        count = int(pinned[0]) & 0xFFFFFFFF
        outs, rows = {}, []
        for group in opt.param_groups:
            name = group["name"]
            old = group["params"][0]
            src = src_tensors[name].detach().contiguous()
            if src.shape[0] != P_src or tuple(src.shape[1:]) != tuple(old.shape[1:]) or src.dtype != torch.float32:
                raise RuntimeError(f"source tensor {name!r} does not match the destination rows")
            P_dst = old.shape[0]
            new = torch.empty((P_dst + count,) + tuple(old.shape[1:]), dtype=old.dtype, device=dev)
            new[:P_dst].copy_(old.detach())
            outs[name] = new
            if count and src.numel():
                row_words = src.numel() // P_src
                rows.append(nat.RowTensor(src.data_ptr(), new.data_ptr() + 4 * P_dst * row_words, row_words, 0))
                rows[-1]._keepalive = src
        if rows:
            arr = (nat.RowTensor * len(rows))(*rows)
            nat.check(lib.htgs_rows_compact(P_src, keep.data_ptr(), scratch.data_ptr(), len(rows), arr, st), "htgs_rows_compact")
        new_params = {}
        for group in opt.param_groups:
            name = group["name"]
            old = group["params"][0]
            stored = opt.state.get(old, None)
            newp = torch.nn.Parameter(outs[name].requires_grad_(True))
            if stored is not None:
                for key in ("exp_avg", "exp_avg_sq"):
                    if key in stored:
                        ext = torch.zeros((old.shape[0] + count,) + tuple(old.shape[1:]), dtype=stored[key].dtype, device=dev)
                        ext[:old.shape[0]].copy_(stored[key])
                        stored[key] = ext
                del opt.state[old]
                opt.state[newp] = stored
            group["params"][0] = newp
            new_params[name] = newp
    _set_model_params(model, new_params)
    n = model._xyz.shape[0]
    model.xyz_gradient_accum = torch.zeros((n, 1), device=dev)
    model.denom = torch.zeros((n, 1), device=dev)
    model.max_radii2D = torch.zeros((n,), device=dev)
    return count


def merge_two_models(dst, src, transform_matrix, mask_dst, mask_src):
    """The tensor work of merge_two_3DGS (trainer/ht3dgs_trainer.py:240-270): prune the destination model by `mask_dst`,
    carry the source positions into the destination frame, drop the source rows in `mask_src`, append the rest.
    `dst` / `src` are HTGaussianModel-like objects; `dst` keeps its optimizer (state extended with zero rows)."""
    prune_model(dst, mask_dst)
    aligned = transform_points(src._xyz, transform_matrix)
    tensors = {"xyz": aligned, "f_dc": src._features_dc, "f_rest": src._features_rest, "opacity": src._opacity,
               "scaling": src._scaling, "rotation": src._rotation}
    return append_rows(dst, tensors, ~mask_src)


# ---- checkpoint tuple ----

def use_fused_adam(model):
    """Swaps the model's torch.optim.Adam (training_setup, gaussian_model_ht.py:289) for this library's FusedAdam, keeping
    parameter groups (names, learning rates) and state; returns the new optimizer."""
    old = model.optimizer
    if isinstance(old, FusedAdam):
        return old
    d = old.defaults
    groups = [{k: v for k, v in g.items() if k in ("params", "lr", "name")} for g in old.param_groups]
    new = FusedAdam(groups, lr=d.get("lr", 0.0), betas=d.get("betas", (0.9, 0.999)), eps=d.get("eps", 1e-15))
    for g in old.param_groups:
        for p in g["params"]:
            if p in old.state:
                new.state[p] = old.state[p]
    model.optimizer = new
    return new


def capture(model):
    """HTGaussianModel.capture (gaussian_model_ht.py:90-105): the 13-field tuple, optimizer state as `state_dict()`
    (FusedAdam keeps torch.optim.Adam's state layout, so the tuple is interchangeable with the reference's)."""
    return (model.active_sh_degree, model._xyz, model._features_dc, model._features_rest, model._scaling, model._rotation,
            model._opacity, model.max_radii2D, model.xyz_gradient_accum, model.denom, model.optimizer.state_dict(),
            model.spatial_lr_scale, getattr(model, "P", None))


def restore(model, model_args, training_args, fused=True):
    """HTGaussianModel.restore (gaussian_model_ht.py:107-124): fields back in place, `training_setup` rebuilds the six
    parameter groups, optimizer state loaded; with fused=True the rebuilt optimizer is replaced by FusedAdam first."""
    (model.active_sh_degree, model._xyz, model._features_dc, model._features_rest, model._scaling, model._rotation,
     model._opacity, model.max_radii2D, xyz_gradient_accum, denom, opt_dict, model.spatial_lr_scale, model.P) = model_args
    model.training_setup(training_args)
    model.xyz_gradient_accum = xyz_gradient_accum
    model.denom = denom
    if fused:
        use_fused_adam(model)
    model.optimizer.load_state_dict(opt_dict)
    return model
