"""CPU oracle of the per-step tail.  TEST INFRASTRUCTURE ONLY.

The optimizer is a third-party dependency of the reference: torch.optim.Adam, constructed at
scene/gaussian_model_ht.py:289 as Adam(l, lr=0.0, eps=1e-15) with default betas (requirements.txt pins
no torch version; this container has 2.11).  Its published algorithm (Kingma & Ba 2015, as implemented by
torch/optim/adam.py `_single_tensor_adam`, non-amsgrad, no weight decay) is restated here in numpy fp32;
tests pin this restatement against torch.optim.Adam itself, which is importable on the CPU.
Densification statistics restate trainer/ht3dgs_trainer.py:143-148 and scene/gaussian_model_ht.py:718-721.
"""
import numpy as np


def adam_step(param, grad, exp_avg, exp_avg_sq, step, lr, beta1=0.9, beta2=0.999, eps=1e-8):
    """One Adam step on fp32 numpy arrays, in place; `step` counts from 1."""
    f = np.float32
    exp_avg += f(1.0 - beta1) * (grad - exp_avg)                             # lerp_(grad, 1 - beta1): weight formed in fp64
    exp_avg_sq *= f(beta2)
    exp_avg_sq += f(1.0 - beta2) * grad * grad                               # addcmul_(grad, grad, value=1 - beta2)
    bc1 = 1.0 - beta1 ** step
    bc2 = 1.0 - beta2 ** step
    step_size = f(lr / bc1)
    denom = np.sqrt(exp_avg_sq) / f(np.sqrt(bc2)) + f(eps)
    param -= step_size * (exp_avg / denom)                                   # addcdiv_(exp_avg, denom, -step_size)


def densification_stats(radii, viewspace_grad, max_radii2D, xyz_gradient_accum, denom):
    """In place; visibility_filter = radii > 0 (scene/gaussian_model_ht.py:892)."""
    vis = radii > 0
    max_radii2D[vis] = np.maximum(max_radii2D[vis], radii[vis].astype(np.float32))
    xyz_gradient_accum[vis] += np.sqrt(viewspace_grad[vis, 0] ** 2 + viewspace_grad[vis, 1] ** 2).astype(np.float32)
    denom[vis] += np.float32(1.0)
