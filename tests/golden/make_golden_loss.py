"""Generates tests/golden/loss_vectors.npz by importing the reference's own loss code
(trainer/losses.py: l1_loss, SSIM_V2) from /root/reference and differentiating it with torch autograd on
the CPU.  Run in the build container:  python tests/golden/make_golden_loss.py"""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, "/root/reference")
from trainer.losses import SSIM_V2, l1_loss  # noqa: E402


def main():
    out = {}
    ssim = SSIM_V2()
    for name, (C, H, W), seed in (("a", (3, 37, 50), 1), ("b", (3, 64, 48), 2), ("c", (1, 9, 7), 3)):
        g = torch.Generator().manual_seed(seed)
        gt = torch.rand(C, H, W, generator=g)
        pred = (gt + 0.2 * torch.randn(C, H, W, generator=g)).clamp(0, 1)
        pred[0, :3, :3] = gt[0, :3, :3]  # exact ties: sign(0) = 0 in the L1 term
        pred.requires_grad_(True)
        lam = 0.2
        l1 = l1_loss(pred, gt)
        dssim = 1 - ssim(pred, gt)
        loss = (1 - lam) * l1 + lam * dssim  # trainer/losses.py:107,123 with lambda_depth = 0
        loss.backward()
        out[f"{name}_pred"] = pred.detach().numpy()
        out[f"{name}_gt"] = gt.numpy()
        out[f"{name}_vals"] = np.array([loss.item(), l1.item(), dssim.item()], np.float64)
        out[f"{name}_grad"] = pred.grad.numpy()
    np.savez_compressed(os.path.join(HERE, "loss_vectors.npz"), **out)
    print({k: v.shape for k, v in out.items()})


if __name__ == "__main__":
    main()
