"""Row surgery kernels (htgs_row_scan / htgs_rows_compact / htgs_rows_append) against the
torch expressions they replace in scene/gaussian_model_ht.py: `t[mask]` (_prune_optimizer :548-566, prune_points
:568-583) and `torch.cat((t, t[mask]))` / `torch.cat((state, zeros))` (cat_tensors_to_optimizer :585-611 as used by
densify_and_clone :656-672).  Pure copies, so the comparison is bit-exact."""
import pytest
import torch

pytestmark = pytest.mark.gpu

SHAPES = {"xyz": (3,), "f_dc": (1, 3), "f_rest": (15, 3), "opacity": (1,), "scaling": (3,), "rotation": (4,)}


def _tensors(P, seed):
    g = torch.Generator().manual_seed(seed)
    return {k: torch.randn((P,) + s, generator=g).cuda() for k, s in SHAPES.items()}


@pytest.mark.parametrize("P,frac", [(1, 1.0), (1, 0.0), (255, 0.5), (256, 0.5), (257, 0.03), (100_003, 0.9), (262_144, 0.5),
                                    (300_000, 0.0), (300_000, 1.0)])
def test_select_and_append_match_torch_indexing(P, frac):
    import htgs_optim
    t = _tensors(P, 1)
    g = torch.Generator().manual_seed(P)
    mask = (torch.rand(P, generator=g) < frac).cuda()
    extra_i = torch.randint(0, 1 << 30, (P,), generator=g, dtype=torch.int32).cuda()   # any 4-byte dtype
    outs = htgs_optim.select_rows(mask, list(t.values()) + [extra_i])
    for o, src in zip(outs, list(t.values()) + [extra_i]):
        assert o.dtype == src.dtype and torch.equal(o, src[mask])
    zero_new = [False, True, False, True, False, True]
    outs = htgs_optim.append_selected_rows(mask, list(t.values()), zero_new)
    for o, src, z in zip(outs, t.values(), zero_new):
        tail = torch.zeros_like(src[mask]) if z else src[mask]
        assert torch.equal(o, torch.cat((src, tail), dim=0))
    # uint8 masks (what `prune_filter` becomes after torch.cat with a bool zeros tensor) behave the same
    outs8 = htgs_optim.select_rows(mask.to(torch.uint8), [t["f_rest"]])
    assert torch.equal(outs8[0], t["f_rest"][mask])


def _make_optimizer(P, cls):
    import htgs_optim
    t = _tensors(P, 7)
    lrs = dict(xyz=1.6e-4, f_dc=2.5e-3, f_rest=1.25e-4, opacity=5e-2, scaling=5e-3, rotation=1e-3)
    params = {k: torch.nn.Parameter(v.clone().requires_grad_(True)) for k, v in t.items()}
    opt = cls([{"params": [params[k]], "lr": lrs[k], "name": k} for k in params], lr=0.0, eps=1e-15)
    g = torch.Generator().manual_seed(11)
    for _ in range(2):  # populate exp_avg / exp_avg_sq / step
        for p in params.values():
            p.grad = torch.randn(p.shape, generator=g).cuda()
        opt.step()
    return opt


def _reference_prune(opt, mask):  # scene/gaussian_model_ht.py:548-566, verbatim semantics
    out = {}
    for group in opt.param_groups:
        stored = opt.state.get(group["params"][0], None)
        if stored is not None:
            stored["exp_avg"] = stored["exp_avg"][mask]
            stored["exp_avg_sq"] = stored["exp_avg_sq"][mask]
            del opt.state[group["params"][0]]
            group["params"][0] = torch.nn.Parameter(group["params"][0][mask].requires_grad_(True))
            opt.state[group["params"][0]] = stored
        else:
            group["params"][0] = torch.nn.Parameter(group["params"][0][mask].requires_grad_(True))
        out[group["name"]] = group["params"][0]
    return out


def _reference_clone(opt, mask):  # densify_and_clone + cat_tensors_to_optimizer
    out = {}
    for group in opt.param_groups:
        ext = group["params"][0][mask]
        stored = opt.state.get(group["params"][0], None)
        stored["exp_avg"] = torch.cat((stored["exp_avg"], torch.zeros_like(ext)), dim=0)
        stored["exp_avg_sq"] = torch.cat((stored["exp_avg_sq"], torch.zeros_like(ext)), dim=0)
        del opt.state[group["params"][0]]
        group["params"][0] = torch.nn.Parameter(torch.cat((group["params"][0], ext), dim=0).requires_grad_(True))
        opt.state[group["params"][0]] = stored
        out[group["name"]] = group["params"][0]
    return out


@pytest.mark.parametrize("op", ["prune", "clone"])
def test_optimizer_surgery_matches_reference_and_training_continues(op):
    import htgs_optim
    P = 20_001
    mask = (torch.rand(P, generator=torch.Generator().manual_seed(3)) < (0.8 if op == "prune" else 0.1)).cuda()
    a = _make_optimizer(P, htgs_optim.FusedAdam)
    b = _make_optimizer(P, htgs_optim.FusedAdam)
    stats = [torch.rand(P, 1).cuda(), torch.rand(P, 1).cuda(), torch.rand(P).cuda()]
    if op == "prune":
        new_a, extra = htgs_optim.prune_optimizer(a, mask, extra=stats)
        new_b = _reference_prune(b, mask)
        for e, s in zip(extra, stats):
            assert torch.equal(e, s[mask])
    else:
        new_a = htgs_optim.clone_selected_to_optimizer(a, mask)
        new_b = _reference_clone(b, mask)
    assert set(new_a) == set(new_b)
    for k in new_a:
        assert isinstance(new_a[k], torch.nn.Parameter) and new_a[k].requires_grad
        assert torch.equal(new_a[k], new_b[k]), k
        sa, sb = a.state[new_a[k]], b.state[new_b[k]]
        assert torch.equal(sa["exp_avg"], sb["exp_avg"]) and torch.equal(sa["exp_avg_sq"], sb["exp_avg_sq"]), k
        assert float(sa["step"]) == float(sb["step"]) == 2.0
    # both optimizers keep stepping identically on the new rows
    g = torch.Generator().manual_seed(5)
    for k in new_a:
        gr = torch.randn(new_a[k].shape, generator=g).cuda()
        new_a[k].grad, new_b[k].grad = gr, gr.clone()
    a.step(); b.step()
    for k in new_a:
        assert torch.equal(new_a[k], new_b[k]), k


def _reference_split(opt, sel, N, new_xyz, new_scaling):
    """densify_and_split (:630-654): densification_postfix(new rows) then prune_points(cat(sel, zeros))."""
    new = {}
    for group in opt.param_groups:
        p = group["params"][0]
        new[group["name"]] = {"xyz": new_xyz, "scaling": new_scaling}.get(group["name"])
        if new[group["name"]] is None:
            new[group["name"]] = p[sel].repeat(N, *([1] * (p.dim() - 1)))
    for group in opt.param_groups:  # cat_tensors_to_optimizer
        ext = new[group["name"]]
        stored = opt.state.get(group["params"][0], None)
        stored["exp_avg"] = torch.cat((stored["exp_avg"], torch.zeros_like(ext)), dim=0)
        stored["exp_avg_sq"] = torch.cat((stored["exp_avg_sq"], torch.zeros_like(ext)), dim=0)
        del opt.state[group["params"][0]]
        group["params"][0] = torch.nn.Parameter(torch.cat((group["params"][0], ext), dim=0).requires_grad_(True))
        opt.state[group["params"][0]] = stored
    prune_filter = torch.cat((sel, torch.zeros(N * int(sel.sum()), device="cuda", dtype=bool)))
    return _reference_prune(opt, ~prune_filter)


@pytest.mark.parametrize("N,frac", [(2, 0.07), (3, 0.5), (2, 0.0), (2, 1.0)])
def test_split_matches_postfix_then_prune(N, frac):
    import htgs_optim
    P = 10_007
    sel = (torch.rand(P, generator=torch.Generator().manual_seed(N)) < frac).cuda()
    count = int(sel.sum())
    a = _make_optimizer(P, htgs_optim.FusedAdam)
    b = _make_optimizer(P, htgs_optim.FusedAdam)
    g = torch.Generator().manual_seed(99)
    new_xyz, new_scaling = torch.randn(N * count, 3, generator=g).cuda(), torch.randn(N * count, 3, generator=g).cuda()
    new_a = htgs_optim.split_selected_in_optimizer(a, sel, N, {"xyz": new_xyz, "scaling": new_scaling})
    new_b = _reference_split(b, sel, N, new_xyz, new_scaling)
    for k in new_b:
        assert new_a[k].shape[0] == P - count + N * count
        assert torch.equal(new_a[k], new_b[k]), k
        sa, sb = a.state[new_a[k]], b.state[new_b[k]]
        assert torch.equal(sa["exp_avg"], sb["exp_avg"]) and torch.equal(sa["exp_avg_sq"], sb["exp_avg_sq"]), k
    with pytest.raises(RuntimeError):
        htgs_optim.split_selected_in_optimizer(a, sel[:-1], N, {})


def test_row_ops_reject_bad_arguments():
    import htgs_optim
    t = torch.zeros(10, 3).cuda()
    with pytest.raises(RuntimeError):
        htgs_optim.select_rows(torch.ones(10, dtype=torch.bool), [t])            # host mask
    with pytest.raises(RuntimeError):
        htgs_optim.select_rows(torch.ones(9, dtype=torch.bool).cuda(), [t])      # row count mismatch
    with pytest.raises(RuntimeError):
        htgs_optim.select_rows(torch.ones(10, dtype=torch.bool).cuda(), [t.double()])  # 8-byte elements
    with pytest.raises(RuntimeError):
        htgs_optim.select_rows(torch.ones(10).cuda(), [t])                       # float mask
