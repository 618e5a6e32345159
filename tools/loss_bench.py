"""Times the fused L1+SSIM loss (fwd+bwd) at 3x1080x1920 against the reference formula run by torch
(grouped conv2d + autograd, what trainer/losses.py does).  Prints one JSON object."""
import importlib, json, os, sys
import torch
import torch.nn.functional as F
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
importlib.import_module("3dgs_hierarchical_training_b200").install()
import htgs_loss
from oracle import loss_ssim


def timeit(fn, warm=5, it=30):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(it):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / it


def main():
    out = {}
    for name, (C, H, W) in (("1080p", (3, 1080, 1920)), ("540p", (3, 540, 960))):
        gt = torch.rand(C, H, W, device="cuda")
        pred = (gt + 0.1 * torch.randn_like(gt)).clamp(0, 1).requires_grad_(True)
        w1 = torch.from_numpy(loss_ssim.window()).float().cuda()
        win = (w1[:, None] @ w1[None, :]).expand(C, 1, 11, 11).contiguous()

        def ref():
            pred.grad = None
            x, y = pred[None], gt[None]
            mu1, mu2 = F.conv2d(x, win, padding=5, groups=C), F.conv2d(y, win, padding=5, groups=C)
            s1 = F.conv2d(x * x, win, padding=5, groups=C) - mu1.pow(2)
            s2 = F.conv2d(y * y, win, padding=5, groups=C) - mu2.pow(2)
            s12 = F.conv2d(x * y, win, padding=5, groups=C) - mu1 * mu2
            m = ((2 * mu1 * mu2 + 1e-4) * (2 * s12 + 9e-4)) / ((mu1.pow(2) + mu2.pow(2) + 1e-4) * (s1 + s2 + 9e-4))
            (0.8 * torch.abs(pred - gt).mean() + 0.2 * (1 - m.mean())).backward()

        def mine():
            pred.grad = None
            loss, _ = htgs_loss.fused_l1_ssim(pred, gt, 0.2)
            loss.backward()

        out[name] = {"torch_reference_formula_ms": timeit(ref), "htgs_fused_ms": timeit(mine),
                     "algorithmic_bytes": 8 * C * H * W * 4}
        out[name]["htgs_GBps"] = out[name]["algorithmic_bytes"] / out[name]["htgs_fused_ms"] / 1e6
    print(json.dumps(out))


if __name__ == "__main__":
    main()
