"""Fused L1 + SSIM loss.  Golden vectors come from the reference's own
trainer/losses.py (SSIM_V2, l1_loss) differentiated by torch autograd on the CPU.  Tolerances: loss value
1e-6 absolute, gradient 1e-5 relative to its largest element (fp32 kernels vs the reference's fp32 conv2d)."""
import os

import numpy as np
import pytest
import torch

from oracle import loss_ssim as oracle

GOLD = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "loss_vectors.npz"))
CASES = ["a", "b", "c"]


@pytest.mark.parametrize("name", CASES)
def test_oracle_matches_reference_loss(name):
    loss, l1, dssim, grad = oracle.l1_ssim(GOLD[f"{name}_pred"], GOLD[f"{name}_gt"], 0.2)
    ref = GOLD[f"{name}_vals"]
    assert abs(loss - ref[0]) <= 1e-6 and abs(l1 - ref[1]) <= 1e-6 and abs(dssim - ref[2]) <= 1e-6
    rg = GOLD[f"{name}_grad"]
    assert np.abs(grad - rg).max() <= 1e-5 * np.abs(rg).max()


@pytest.mark.gpu
@pytest.mark.parametrize("name", CASES)
def test_cuda_loss_matches_golden_and_oracle(name):
    import htgs_loss
    pred = torch.from_numpy(GOLD[f"{name}_pred"]).cuda().requires_grad_(True)
    gt = torch.from_numpy(GOLD[f"{name}_gt"]).cuda()
    loss, parts = htgs_loss.fused_l1_ssim(pred, gt, 0.2)
    (3.0 * loss).backward()  # a non-unit upstream gradient
    ref = GOLD[f"{name}_vals"]
    assert np.abs(parts.cpu().numpy().astype(np.float64) - ref).max() <= 1e-6
    rg = GOLD[f"{name}_grad"]
    assert np.abs(pred.grad.cpu().numpy() / 3.0 - rg).max() <= 1e-5 * np.abs(rg).max()
    o = oracle.l1_ssim(GOLD[f"{name}_pred"], GOLD[f"{name}_gt"], 0.2)
    assert abs(float(loss.detach()) - o[0]) <= 1e-6


@pytest.mark.gpu
def test_cuda_loss_full_size_against_torch_formula():
    """1080p: compare with the reference formula evaluated by torch on the GPU (conv2d + autograd)."""
    import torch.nn.functional as F
    import htgs_loss
    g = torch.Generator().manual_seed(7)
    C, H, W = 3, 1080, 1920
    gt = torch.rand(C, H, W, generator=g).cuda()
    pred0 = (gt + 0.1 * torch.randn(C, H, W, generator=g).cuda()).clamp(0, 1)
    w1 = torch.from_numpy(oracle.window()).float().cuda()
    win = (w1[:, None] @ w1[None, :]).expand(C, 1, 11, 11).contiguous()
    a = pred0.clone().requires_grad_(True)
    x, y = a[None], gt[None]
    mu1, mu2 = F.conv2d(x, win, padding=5, groups=C), F.conv2d(y, win, padding=5, groups=C)
    s1 = F.conv2d(x * x, win, padding=5, groups=C) - mu1 * mu1
    s2 = F.conv2d(y * y, win, padding=5, groups=C) - mu2 * mu2
    s12 = F.conv2d(x * y, win, padding=5, groups=C) - mu1 * mu2
    m = ((2 * mu1 * mu2 + 1e-4) * (2 * s12 + 9e-4)) / ((mu1 * mu1 + mu2 * mu2 + 1e-4) * (s1 + s2 + 9e-4))
    ref = 0.8 * (a - gt).abs().mean() + 0.2 * (1 - m.mean())
    ref.backward()
    b = pred0.clone().requires_grad_(True)
    loss, parts = htgs_loss.fused_l1_ssim(b, gt, 0.2)
    loss.backward()
    assert abs(float(loss) - float(ref)) <= 2e-6
    assert float((b.grad - a.grad).abs().max()) <= 1e-5 * float(a.grad.abs().max())
    with torch.no_grad():  # no derivative maps are produced when no gradient is needed
        l2, _ = htgs_loss.fused_l1_ssim(pred0, gt, 0.2)
    assert float(l2) == float(loss)
    with pytest.raises(RuntimeError):
        htgs_loss.fused_l1_ssim(pred0.cpu(), gt.cpu(), 0.2)
