"""Opt-in fused image loss: the loss HTGaussianTrainer.compute_loss builds from
trainer/losses.py -- (1 - lambda_dssim) * L1 + lambda_dssim * (1 - SSIM) -- as one autograd Function backed by
two CUDA kernels (htgs_l1_ssim_forward / _backward).  Its gradient is exactly the dL_dcolor the rasterizer
backward consumes.  CUDA only; no fallback."""
import ctypes

import torch

import htgs_native as nat


def _stream(dev):
    return ctypes.c_void_p(torch.cuda.current_stream(dev).cuda_stream)


class _L1SSIM(torch.autograd.Function):
    @staticmethod
    def forward(ctx, pred, gt, lambda_dssim):
        if not pred.is_cuda or not gt.is_cuda:
            raise RuntimeError("fused_l1_ssim: CUDA tensors required (no CPU path)")
        if pred.dim() != 3 or pred.shape != gt.shape or pred.dtype != torch.float32 or gt.dtype != torch.float32:
            raise RuntimeError("fused_l1_ssim: pred and gt must be float32 [C,H,W] of the same shape")
        pred_c, gt_c = pred.contiguous(), gt.contiguous()
        C, H, W = (int(v) for v in pred.shape)
        dev = pred.device
        lib = nat.lib()
        need_grad = ctx.needs_input_grad[0]
        with torch.cuda.device(dev):
            out3 = torch.empty(3, dtype=torch.float32, device=dev)
            scratch = torch.empty(lib.htgs_l1_ssim_scratch_bytes(C, H, W), dtype=torch.uint8, device=dev)
            dmaps = torch.empty((3, C, H, W), dtype=torch.float32, device=dev) if need_grad else None
            nat.check(lib.htgs_l1_ssim_forward(C, H, W, pred_c.data_ptr(), gt_c.data_ptr(), float(lambda_dssim), nat.ptr(dmaps),
                                               scratch.data_ptr(), out3.data_ptr(), _stream(dev)), "htgs_l1_ssim_forward")
        ctx.lambda_dssim = float(lambda_dssim)
        if need_grad:
            ctx.save_for_backward(pred_c, gt_c, dmaps)
        ctx.mark_non_differentiable(out3)
        return out3[0], out3

    @staticmethod
    def backward(ctx, grad_loss, _grad_parts):
        pred, gt, dmaps = ctx.saved_tensors
        C, H, W = (int(v) for v in pred.shape)
        dev = pred.device
        g = grad_loss.to(device=dev, dtype=torch.float32).reshape(1).contiguous()
        with torch.cuda.device(dev):
            dpred = torch.empty_like(pred)
            nat.check(nat.lib().htgs_l1_ssim_backward(C, H, W, pred.data_ptr(), gt.data_ptr(), dmaps.data_ptr(), ctx.lambda_dssim,
                                                      g.data_ptr(), dpred.data_ptr(), _stream(dev)), "htgs_l1_ssim_backward")
        return dpred, None, None


def fused_l1_ssim(rgb_pred, rgb_gt, lambda_dssim=0.2):
    """Returns (loss, parts) with parts = tensor([loss, l1, 1 - ssim]); differentiable w.r.t. rgb_pred.
    Same value as trainer/losses.py Loss.forward's `rgb_full_loss + lambda_dssim * dssim_loss`."""
    return _L1SSIM.apply(rgb_pred, rgb_gt, lambda_dssim)
