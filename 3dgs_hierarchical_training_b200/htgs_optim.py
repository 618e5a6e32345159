"""Opt-in fused versions of the per-step tail of HTGaussianTrainer.train_step.

`FusedAdam` is a torch.optim.Optimizer with the state layout of torch.optim.Adam (`state[p]["step"]`,
`["exp_avg"]`, `["exp_avg_sq"]`, `param_groups[i]["lr"/"name"/"params"]`), so the model code that edits
optimizer state when it densifies or prunes (scene/gaussian_model_ht.py `cat_tensors_to_optimizer`,
`_prune_optimizer`, `replace_tensor_to_optimizer`) keeps working; only `step()` differs: one CUDA launch for
all groups instead of torch's per-op foreach kernels.  Construct it exactly like the reference does:

    optimizer = FusedAdam(l, lr=0.0, eps=1e-15)         # scene/gaussian_model_ht.py:289

`add_densification_stats` replaces ht3dgs_trainer.py:143-148 + gaussian_model_ht.py:718-721 without the
host synchronisations of boolean-mask indexing.  CUDA only; no fallback.
"""
import ctypes

import torch

import htgs_native as nat


def _stream(dev):
    return ctypes.c_void_p(torch.cuda.current_stream(dev).cuda_stream)


def add_densification_stats(radii, viewspace_grad, max_radii2D, xyz_gradient_accum, denom):
    """In place, for radii > 0: max_radii2D = max(max_radii2D, radii); xyz_gradient_accum += ||grad[:, :2]||;
    denom += 1.  `viewspace_grad` is `viewspace_point_tensor.grad` ([P,3] float32)."""
    for t, name in ((radii, "radii"), (viewspace_grad, "viewspace_grad"), (max_radii2D, "max_radii2D"),
                    (xyz_gradient_accum, "xyz_gradient_accum"), (denom, "denom")):
        if not t.is_cuda or not t.is_contiguous():
            raise RuntimeError(f"add_densification_stats: {name} must be a contiguous CUDA tensor")
    if radii.dtype != torch.int32 or any(t.dtype != torch.float32 for t in (viewspace_grad, max_radii2D, xyz_gradient_accum, denom)):
        raise RuntimeError("add_densification_stats: radii int32, everything else float32")
    P = int(radii.numel())
    if viewspace_grad.shape != (P, 3) or max_radii2D.numel() != P or xyz_gradient_accum.numel() != P or denom.numel() != P:
        raise RuntimeError("add_densification_stats: shape mismatch")
    dev = radii.device
    with torch.cuda.device(dev):
        nat.check(nat.lib().htgs_densification_stats(P, radii.data_ptr(), viewspace_grad.data_ptr(), max_radii2D.data_ptr(),
                                                     xyz_gradient_accum.data_ptr(), denom.data_ptr(), _stream(dev)),
                  "htgs_densification_stats")


class FusedAdam(torch.optim.Optimizer):
    def __init__(self, params, lr=1e-3, betas=(0.9, 0.999), eps=1e-8):
        if not 0.0 <= lr or not 0.0 <= eps or not 0.0 <= betas[0] < 1.0 or not 0.0 <= betas[1] < 1.0:
            raise ValueError("invalid Adam hyper-parameter")
        super().__init__(params, dict(lr=lr, betas=betas, eps=eps))

    @torch.no_grad()
    def step(self, closure=None):
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        # groups sharing (betas, eps, step) go into one launch, at most HTGS_ADAM_MAX_GROUPS tensors each
        batches = {}
        for group in self.param_groups:
            for p in group["params"]:
                if p.grad is None:
                    continue
                if not p.is_cuda or p.dtype != torch.float32 or not p.is_contiguous():
                    raise RuntimeError("FusedAdam: parameters must be contiguous float32 CUDA tensors (no CPU path)")
                if p.grad.is_sparse:
                    raise RuntimeError("FusedAdam does not support sparse gradients")
                st = self.state[p]
                if len(st) == 0:
                    st["step"] = torch.tensor(0.0)
                    st["exp_avg"] = torch.zeros_like(p, memory_format=torch.preserve_format)
                    st["exp_avg_sq"] = torch.zeros_like(p, memory_format=torch.preserve_format)
                st["step"] += 1
                key = (p.device, group["betas"], group["eps"], int(st["step"]))
                batches.setdefault(key, []).append((p, p.grad.contiguous(), st["exp_avg"], st["exp_avg_sq"], float(group["lr"])))
        lib = nat.lib()
        for (dev, betas, eps, step), items in batches.items():
            with torch.cuda.device(dev):
                for i in range(0, len(items), 8):
                    chunk = items[i:i + 8]
                    arr = (nat.AdamGroup * len(chunk))()
                    for k, (p, g, m, v, lr) in enumerate(chunk):
                        arr[k] = nat.AdamGroup(p.data_ptr(), g.data_ptr(), m.data_ptr(), v.data_ptr(), p.numel(), lr)
                    nat.check(lib.htgs_adam_step(len(chunk), arr, float(betas[0]), float(betas[1]), float(eps), step, _stream(dev)),
                              "htgs_adam_step")
        return loss
