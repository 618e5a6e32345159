"""CPU oracle of the image loss.  TEST INFRASTRUCTURE ONLY.

numpy restatement of trainer/losses.py: l1_loss :139-140, gaussian/create_window :147-158, _ssim :164-200
(zero-padded 11x11 sigma-1.5 window per channel, C1 = 0.01^2, C2 = 0.03^2, mean over all elements) and of
Loss.forward's combination :107-124 with lambda_depth = 0.  The gradient is derived by hand (float64); tests pin
both value and gradient against the reference's own SSIM_V2 / l1_loss run through torch autograd on the CPU
(tests/golden/make_golden_loss.py)."""
import math

import numpy as np


def window():
    g = np.array([math.exp(-(x - 5) ** 2 / float(2 * 1.5 ** 2)) for x in range(11)], np.float32)
    return (g / g.sum()).astype(np.float64)


def _conv(x, w):
    """'same' correlation with zero padding, separable, x: [C,H,W] float64."""
    C, H, W = x.shape
    p = np.pad(x, ((0, 0), (5, 5), (5, 5)))
    h = sum(w[k] * p[:, :, k:k + W] for k in range(11))
    return sum(w[k] * h[:, k:k + H, :] for k in range(11))


def l1_ssim(pred, gt, lambda_dssim=0.2):
    """Returns (loss, l1, dssim, dloss_dpred) in float64."""
    a, b = pred.astype(np.float64), gt.astype(np.float64)
    w = window()
    C1, C2 = 0.01 ** 2, 0.03 ** 2
    mu1, mu2 = _conv(a, w), _conv(b, w)
    e11, e22, e12 = _conv(a * a, w), _conv(b * b, w), _conv(a * b, w)
    s1, s2, s12 = e11 - mu1 * mu1, e22 - mu2 * mu2, e12 - mu1 * mu2
    A, B, Cd, D = 2 * mu1 * mu2 + C1, 2 * s12 + C2, mu1 * mu1 + mu2 * mu2 + C1, s1 + s2 + C2
    m = A * B / (Cd * D)
    n = a.size
    l1 = np.abs(a - b).mean()
    dssim = 1.0 - m.mean()
    loss = (1 - lambda_dssim) * l1 + lambda_dssim * dssim
    dm_dmu1 = 2 * mu2 * B / (Cd * D) - 2 * mu1 * m / Cd + 2 * mu1 * m / D - 2 * mu2 * A / (Cd * D)
    dm_de11 = -m / D
    dm_de12 = 2 * A / (Cd * D)
    dmap = _conv(dm_dmu1, w) + 2 * a * _conv(dm_de11, w) + b * _conv(dm_de12, w)
    grad = ((1 - lambda_dssim) * np.sign(a - b) - lambda_dssim * dmap) / n
    return loss, l1, dssim, grad
