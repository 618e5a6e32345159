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


# ---- row surgery for prune / clone (first half) ----

def _move_rows(mask, tensors, append, zero_new=None):
    """Shared driver: one mask scan, one host read of the selected count, one launch per <= 24 tensors."""
    if mask.dtype not in (torch.bool, torch.uint8) or not mask.is_cuda or mask.dim() != 1:
        raise RuntimeError("row mask must be a 1-D bool/uint8 CUDA tensor")
    P = int(mask.numel())
    dev = mask.device
    mask = mask.contiguous()
    srcs = []
    for t in tensors:
        if not t.is_cuda or t.device != dev:
            raise RuntimeError("row tensors must live on the mask's CUDA device")
        if t.size(0) != P:
            raise RuntimeError(f"row tensor has {t.size(0)} rows, mask has {P}")
        if t.element_size() != 4:
            raise RuntimeError("row tensors must have 4-byte elements")
        srcs.append(t.detach().contiguous())
    lib = nat.lib()
    with torch.cuda.device(dev):
        st = _stream(dev)
        scratch = torch.empty(lib.htgs_row_scan_scratch_bytes(P), dtype=torch.uint8, device=dev)
        pinned = torch.zeros(1, dtype=torch.int32).pin_memory()
        nat.check(lib.htgs_row_scan(P, mask.data_ptr(), scratch.data_ptr(), pinned.data_ptr(), st), "htgs_row_scan")
        torch.cuda.current_stream(dev).synchronize()  # the one host read: output sizes depend on it
        count = int(pinned[0]) & 0xFFFFFFFF
        n_out = P + count if append else count
        outs = [torch.empty((n_out,) + tuple(t.shape[1:]), dtype=t.dtype, device=dev) for t in srcs]
        fn, name = (lib.htgs_rows_append, "htgs_rows_append") if append else (lib.htgs_rows_compact, "htgs_rows_compact")
        for lo in range(0, len(srcs), nat.ROWS_MAX_TENSORS):
            part = [(k, srcs[k]) for k in range(lo, min(lo + nat.ROWS_MAX_TENSORS, len(srcs))) if srcs[k].numel() > 0]
            if not part or P == 0 or n_out == 0:
                continue
            arr = (nat.RowTensor * len(part))()
            for j, (k, t) in enumerate(part):
                arr[j] = nat.RowTensor(t.data_ptr(), outs[k].data_ptr(), t.numel() // P, int(bool(zero_new[k])) if zero_new else 0)
            nat.check(fn(P, mask.data_ptr(), scratch.data_ptr(), len(part), arr, st), name)
    return outs, count


def select_rows(mask, tensors):
    """[t[mask] for t in tensors] with one scan, one sync and one launch (torch: a nonzero + sync + gather each)."""
    return _move_rows(mask, tensors, append=False)[0]


def append_selected_rows(mask, tensors, zero_new=None):
    """[torch.cat((t, t[mask])) for t in tensors]; where zero_new[k] is true the appended rows are zeros instead."""
    return _move_rows(mask, tensors, append=True, zero_new=zero_new)[0]


def prune_optimizer(optimizer, keep_mask, extra=()):
    """`_prune_optimizer` (scene/gaussian_model_ht.py:548-566) for every group at once, plus `extra` tensors
    (xyz_gradient_accum, denom, max_radii2D of prune_points :568-583) in the same launch.
    Returns ({group name: new nn.Parameter}, [pruned extra tensors])."""
    groups = [g for g in optimizer.param_groups]
    tensors, slots = [], []
    for gi, g in enumerate(groups):
        if len(g["params"]) != 1:
            raise RuntimeError("prune_optimizer expects one parameter per group (the reference's layout)")
        p = g["params"][0]
        tensors.append(p.data)
        slots.append((gi, "param"))
        st = optimizer.state.get(p, None)
        if st is not None and "exp_avg" in st:
            tensors += [st["exp_avg"], st["exp_avg_sq"]]
            slots += [(gi, "exp_avg"), (gi, "exp_avg_sq")]
    n_opt = len(tensors)
    outs = select_rows(keep_mask, tensors + list(extra))
    new_params = {}
    by_group = {}
    for (gi, what), t in zip(slots, outs[:n_opt]):
        by_group.setdefault(gi, {})[what] = t
    for gi, g in enumerate(groups):
        old = g["params"][0]
        stored = optimizer.state.get(old, None)
        newp = torch.nn.Parameter(by_group[gi]["param"].requires_grad_(True))
        if stored is not None:
            if "exp_avg" in by_group[gi]:
                stored["exp_avg"], stored["exp_avg_sq"] = by_group[gi]["exp_avg"], by_group[gi]["exp_avg_sq"]
            del optimizer.state[old]
            optimizer.state[newp] = stored
        g["params"][0] = newp
        new_params[g.get("name", str(gi))] = newp
    return new_params, outs[n_opt:]


def clone_selected_to_optimizer(optimizer, selected_mask):
    """densify_and_clone's tensor work (scene/gaussian_model_ht.py:656-672 + cat_tensors_to_optimizer :585-611):
    every parameter gets its selected rows appended, both Adam moments get as many zero rows.
    Returns {group name: new nn.Parameter}."""
    groups = [g for g in optimizer.param_groups]
    tensors, zero_new, slots = [], [], []
    for gi, g in enumerate(groups):
        if len(g["params"]) != 1:
            raise RuntimeError("clone_selected_to_optimizer expects one parameter per group")
        p = g["params"][0]
        tensors.append(p.data); zero_new.append(False); slots.append((gi, "param"))
        st = optimizer.state.get(p, None)
        if st is not None and "exp_avg" in st:
            tensors += [st["exp_avg"], st["exp_avg_sq"]]
            zero_new += [True, True]
            slots += [(gi, "exp_avg"), (gi, "exp_avg_sq")]
    outs = append_selected_rows(selected_mask, tensors, zero_new)
    by_group = {}
    for (gi, what), t in zip(slots, outs):
        by_group.setdefault(gi, {})[what] = t
    new_params = {}
    for gi, g in enumerate(groups):
        old = g["params"][0]
        stored = optimizer.state.get(old, None)
        newp = torch.nn.Parameter(by_group[gi]["param"].requires_grad_(True))
        if stored is not None:
            if "exp_avg" in by_group[gi]:
                stored["exp_avg"], stored["exp_avg_sq"] = by_group[gi]["exp_avg"], by_group[gi]["exp_avg_sq"]
            del optimizer.state[old]
            optimizer.state[newp] = stored
        g["params"][0] = newp
        new_params[g.get("name", str(gi))] = newp
    return new_params


def split_selected_in_optimizer(optimizer, selected_mask, N, computed):
    """densify_and_split's tensor work (scene/gaussian_model_ht.py:630-654): what `densification_postfix(new_*)`
    followed by `prune_points(cat(selected, zeros(N * count)))` leaves -- every parameter becomes
    cat(param[~selected], new rows) with new rows = the selected rows repeated N times, except for the groups named in
    `computed` ({group name: [N * count, ...] tensor}; the reference computes new positions and scales, :633-643),
    and both Adam moments become cat(moment[~selected], zeros).  One mask scan, one host read, one launch.
    Returns {group name: new nn.Parameter}."""
    if selected_mask.dtype not in (torch.bool, torch.uint8) or not selected_mask.is_cuda or selected_mask.dim() != 1:
        raise RuntimeError("row mask must be a 1-D bool/uint8 CUDA tensor")
    if N < 1:
        raise RuntimeError("N must be >= 1")
    mask = selected_mask.contiguous()
    P, dev = int(mask.numel()), mask.device
    groups = list(optimizer.param_groups)
    lib = nat.lib()
    with torch.cuda.device(dev):
        st = _stream(dev)
        scratch = torch.empty(lib.htgs_row_scan_scratch_bytes(P), dtype=torch.uint8, device=dev)
        pinned = torch.zeros(1, dtype=torch.int32).pin_memory()
        nat.check(lib.htgs_row_scan(P, mask.data_ptr(), scratch.data_ptr(), pinned.data_ptr(), st), "htgs_row_scan")
        torch.cuda.current_stream(dev).synchronize()
        count = int(pinned[0]) & 0xFFFFFFFF
        n_out = P - count + N * count
        items = []  # (group index, what, src, dst, zero_new)
        for gi, g in enumerate(groups):
            if len(g["params"]) != 1:
                raise RuntimeError("split_selected_in_optimizer expects one parameter per group")
            p = g["params"][0]
            if p.size(0) != P or p.element_size() != 4 or p.device != dev:
                raise RuntimeError("parameter rows / dtype / device do not match the mask")
            name = g.get("name", str(gi))
            new_rows = computed.get(name)
            if new_rows is not None and tuple(new_rows.shape) != (N * count,) + tuple(p.shape[1:]):
                raise RuntimeError(f"computed[{name!r}] must have shape {(N * count,) + tuple(p.shape[1:])}, got {tuple(new_rows.shape)}")
            src = p.data.contiguous()
            items.append((gi, "param", src, torch.empty((n_out,) + tuple(p.shape[1:]), dtype=p.dtype, device=dev),
                          2 if new_rows is not None else 0))
            stt = optimizer.state.get(p, None)
            if stt is not None and "exp_avg" in stt:
                for what in ("exp_avg", "exp_avg_sq"):
                    items.append((gi, what, stt[what].contiguous(), torch.empty((n_out,) + tuple(p.shape[1:]), dtype=p.dtype, device=dev), 1))
        live = [it for it in items if it[2].numel() > 0]
        for lo in range(0, len(live), nat.ROWS_MAX_TENSORS):
            part = live[lo:lo + nat.ROWS_MAX_TENSORS]
            arr = (nat.RowTensor * len(part))()
            for j, (_, _, src, dst, z) in enumerate(part):
                arr[j] = nat.RowTensor(src.data_ptr(), dst.data_ptr(), src.numel() // P, z)
            if n_out > 0:
                nat.check(lib.htgs_rows_split(P, int(N), count, mask.data_ptr(), scratch.data_ptr(), len(part), arr, st), "htgs_rows_split")
        by_group = {}
        for gi, what, _, dst, _ in items:
            by_group.setdefault(gi, {})[what] = dst
        new_params = {}
        for gi, g in enumerate(groups):
            name = g.get("name", str(gi))
            if computed.get(name) is not None and count > 0:
                by_group[gi]["param"][P - count:] = computed[name]
            old = g["params"][0]
            stored = optimizer.state.get(old, None)
            newp = torch.nn.Parameter(by_group[gi]["param"].requires_grad_(True))
            if stored is not None:
                if "exp_avg" in by_group[gi]:
                    stored["exp_avg"], stored["exp_avg_sq"] = by_group[gi]["exp_avg"], by_group[gi]["exp_avg_sq"]
                del optimizer.state[old]
                optimizer.state[newp] = stored
            g["params"][0] = newp
            new_params[name] = newp
    return new_params
