"""Per-step tail: densification statistics and the fused multi-group Adam.

CPU: the numpy oracle is pinned against torch.optim.Adam itself (the reference's actual dependency,
scene/gaussian_model_ht.py:289) and against the reference's masked-indexing expressions for the statistics.
GPU: the CUDA entry points against the oracle and against torch.optim.Adam on the device, over several
steps, ragged sizes (vector and tail paths), zero learning rates, and the state surgery the model performs
when it densifies.  Tolerance: 2e-6 relative (fp32 rounding of a differently associated but identical formula)."""
import numpy as np
import pytest
import torch

from oracle import train_tail as oracle

LRS = dict(xyz=1.6e-4, f_dc=2.5e-3, f_rest=2.5e-3 / 20.0, opacity=5e-2, scaling=5e-3, rotation=1e-3)
SHAPES = dict(xyz=(3,), f_dc=(1, 3), f_rest=(15, 3), opacity=(1,), scaling=(3,), rotation=(4,))


def _params(P, seed, device="cpu"):
    g = torch.Generator().manual_seed(seed)
    return {k: torch.randn((P,) + s, generator=g).to(device).requires_grad_(True) for k, s in SHAPES.items()}


def _groups(params, lrs):
    return [{"params": [params[k]], "lr": lrs[k], "name": k} for k in SHAPES]


def _set_grads(params, step, device="cpu"):
    for i, (k, p) in enumerate(params.items()):
        g = torch.Generator().manual_seed(1000 * step + i)
        p.grad = (torch.randn(p.shape, generator=g) * (10.0 ** ((i % 3) - 2))).to(device)


def test_numpy_oracle_matches_torch_adam_cpu():
    P = 257
    params = _params(P, 3)
    opt = torch.optim.Adam(_groups(params, LRS), lr=0.0, eps=1e-15)
    mine = {k: p.detach().numpy().copy() for k, p in params.items()}
    m = {k: np.zeros_like(v) for k, v in mine.items()}
    v = {k: np.zeros_like(vv) for k, vv in mine.items()}
    for step in range(1, 6):
        _set_grads(params, step)
        for k in SHAPES:
            oracle.adam_step(mine[k], params[k].grad.numpy(), m[k], v[k], step, LRS[k], eps=1e-15)
        opt.step()
        for k in SHAPES:
            ref = params[k].detach().numpy()
            assert np.abs(mine[k] - ref).max() <= 2e-6 * max(1.0, np.abs(ref).max()), (k, step)
            rm, rv = opt.state[params[k]]["exp_avg"].numpy(), opt.state[params[k]]["exp_avg_sq"].numpy()
            assert np.abs(m[k] - rm).max() <= 2e-6 * np.abs(rm).max()
            assert np.abs(v[k] - rv).max() <= 2e-6 * np.abs(rv).max()


def test_stats_oracle_matches_reference_expressions():
    g = torch.Generator().manual_seed(9)
    P = 1000
    radii = torch.randint(-1, 30, (P,), generator=g, dtype=torch.int32).clamp_min(0)
    grad = torch.randn(P, 3, generator=g)
    max_r, acc, den = torch.rand(P, generator=g) * 20, torch.rand(P, 1, generator=g), torch.rand(P, 1, generator=g).round()
    o = [t.numpy().copy() for t in (max_r, acc.squeeze(1), den.squeeze(1))]
    oracle.densification_stats(radii.numpy(), grad.numpy(), *o)
    vis = radii > 0  # the reference's own lines (ht3dgs_trainer.py:143-144, gaussian_model_ht.py:718-721)
    max_r[vis] = torch.max(max_r[vis], radii[vis])
    acc[vis] += torch.norm(grad[vis, :2], dim=-1, keepdim=True)
    den[vis] += 1
    assert np.allclose(o[0], max_r.numpy()) and np.allclose(o[1], acc.squeeze(1).numpy(), rtol=1e-6) and np.allclose(o[2], den.squeeze(1).numpy())


@pytest.mark.gpu
@pytest.mark.parametrize("P", [1, 37, 4096, 100_003])
def test_fused_adam_matches_oracle_and_torch(P):
    import htgs_optim
    dev = "cuda"
    a, b = _params(P, 5, dev), _params(P, 5, dev)
    lrs = dict(LRS, rotation=0.0)  # a frozen group (lr 0, like fix_pos / fix_feat)
    fused = htgs_optim.FusedAdam(_groups(a, lrs), lr=0.0, eps=1e-15)
    ref = torch.optim.Adam(_groups(b, lrs), lr=0.0, eps=1e-15)
    o = {k: p.detach().cpu().numpy().copy() for k, p in a.items()}
    m = {k: np.zeros_like(v) for k, v in o.items()}
    v = {k: np.zeros_like(vv) for k, vv in o.items()}
    for step in range(1, 5):
        _set_grads(a, step, dev)
        _set_grads(b, step, dev)
        for k in SHAPES:
            oracle.adam_step(o[k], a[k].grad.cpu().numpy(), m[k], v[k], step, lrs[k], eps=1e-15)
        fused.step()
        ref.step()
        for k in SHAPES:
            got = a[k].detach().cpu().numpy()
            scale = max(1.0, np.abs(o[k]).max())
            assert np.abs(got - o[k]).max() <= 2e-6 * scale, (k, step, "oracle")
            assert np.abs(got - b[k].detach().cpu().numpy()).max() <= 2e-6 * scale, (k, step, "torch")
    assert torch.equal(a["rotation"], _params(P, 5, dev)["rotation"])  # lr 0 leaves the group untouched
    st = fused.state[a["xyz"]]
    assert set(st) == {"step", "exp_avg", "exp_avg_sq"} and float(st["step"]) == 4.0
    fused.zero_grad(set_to_none=True)
    assert a["xyz"].grad is None


@pytest.mark.gpu
def test_fused_adam_survives_state_surgery():
    """cat_tensors_to_optimizer-style growth (scene/gaussian_model_ht.py densification): replace the parameter and
    extend its moments with zeros, then keep stepping."""
    import htgs_optim
    p = torch.nn.Parameter(torch.randn(10, 3, device="cuda"))
    opt = htgs_optim.FusedAdam([{"params": [p], "lr": 1e-2, "name": "xyz"}], lr=0.0, eps=1e-15)
    p.grad = torch.ones_like(p)
    opt.step()
    group = opt.param_groups[0]
    stored = opt.state.get(group["params"][0])
    ext = torch.randn(5, 3, device="cuda")
    stored["exp_avg"] = torch.cat((stored["exp_avg"], torch.zeros_like(ext)), dim=0)
    stored["exp_avg_sq"] = torch.cat((stored["exp_avg_sq"], torch.zeros_like(ext)), dim=0)
    del opt.state[group["params"][0]]
    group["params"][0] = torch.nn.Parameter(torch.cat((group["params"][0], ext), dim=0).requires_grad_(True))
    opt.state[group["params"][0]] = stored
    q = group["params"][0]
    q.grad = torch.ones_like(q)
    before = q.detach().clone()
    opt.step()
    assert q.shape == (15, 3) and float(opt.state[q]["step"]) == 2.0
    assert bool((q.detach() < before).all())


@pytest.mark.gpu
def test_densification_stats_matches_oracle():
    import htgs_optim
    g = torch.Generator().manual_seed(11)
    P = 50_001
    radii = torch.randint(-3, 40, (P,), generator=g, dtype=torch.int32).clamp_min(0)
    grad = torch.randn(P, 3, generator=g)
    max_r, acc, den = torch.rand(P, generator=g) * 30, torch.rand(P, 1, generator=g), torch.rand(P, 1, generator=g).round()
    o = [t.numpy().copy() for t in (max_r, acc.squeeze(1), den.squeeze(1))]
    oracle.densification_stats(radii.numpy(), grad.numpy(), *o)
    d = [t.cuda() for t in (radii, grad, max_r, acc, den)]
    htgs_optim.add_densification_stats(*d)
    assert np.array_equal(d[2].cpu().numpy(), o[0])
    assert np.allclose(d[3].squeeze(1).cpu().numpy(), o[1], rtol=1e-6, atol=0)
    assert np.array_equal(d[4].squeeze(1).cpu().numpy(), o[2])
    with pytest.raises(RuntimeError):
        htgs_optim.add_densification_stats(radii, grad, max_r, acc, den)  # CPU tensors: no fallback


@pytest.mark.gpu
def test_reduce_densification_stats_single_rank_equals_direct_update():
    """Without a process group the data-parallel wrapper must reduce to the plain per-view update."""
    import htgs_dist
    import htgs_optim
    g = torch.Generator().manual_seed(12)
    P = 20_011
    radii = torch.randint(-3, 40, (P,), generator=g, dtype=torch.int32).clamp_min(0).cuda()
    grad = torch.randn(P, 3, generator=g).cuda()
    state = [torch.rand(P, generator=g) * 30, torch.rand(P, 1, generator=g), torch.rand(P, 1, generator=g).round()]
    a = [t.clone().cuda() for t in state]
    b = [t.clone().cuda() for t in state]
    htgs_optim.add_densification_stats(radii, grad, *a)
    htgs_dist.reduce_densification_stats(radii, grad, *b)
    assert torch.equal(a[0], b[0]) and torch.equal(a[2], b[2])
    assert torch.allclose(a[1], b[1], rtol=1e-6, atol=0)
