"""Deterministic synthetic scenes of the BASELINE.json shapes.

No datasets or checkpoints are available offline, so every test and the bench
use these: Gaussians scattered in the frustum of a Tanks&Temples-Francis-shaped
camera (FoVx from arguments/full/Tanks/Francis.yml:46), matrices stored in the
reference's row-vector convention (scene/cameras.py:76-98,
utils/graphics_utils.py:118-137).  Everything is generated on the CPU from a
seeded torch.Generator and moved to `device` afterwards, so a scene is
bit-identical wherever it is built.
"""
import math

import torch

FOVX_FRANCIS = 1.3541787529604106

# name -> generator arguments; "config N" are the rows of BASELINE.json.configs
CONFIGS = {
    "config1": dict(P=10_000, W=256, H=256, s_max=6.0, seed=1, kind="default"),
    "config2": dict(P=300_000, W=960, H=540, s_max=8.0, seed=2, kind="default"),
    "config3": dict(P=1_000_000, W=1920, H=1080, s_max=8.0, seed=3, kind="default"),
    "config4": dict(P=6_000_000, W=1920, H=1080, s_max=4.0, seed=4, kind="default"),
    "config5": dict(P=2_000_000, W=512, H=512, s_max=0.8, seed=5, kind="stress"),
    # small cases for parity tests
    "tiny": dict(P=600, W=80, H=56, s_max=6.0, seed=11, kind="default"),
    "small": dict(P=4_000, W=200, H=136, s_max=6.0, seed=12, kind="default"),
    "ragged": dict(P=3_000, W=250, H=131, s_max=10.0, seed=13, kind="default"),
    "deep": dict(P=20_000, W=64, H=64, s_max=0.8, seed=14, kind="stress"),
}


def make_camera(W, H, fovx=FOVX_FRANCIS, znear=0.01, zfar=100.0, yaw_deg=0.0, pivot_z=4.5, device="cpu"):
    """Camera at the origin looking down +z, optionally yawed about the point (0,0,pivot_z).

    Returns a dict with viewmatrix / projmatrix (= view @ proj, "full_proj_transform") /
    campos / tanfovx / tanfovy exactly as CF3DGS_Render.render feeds them to
    GaussianRasterizationSettings (scene/gaussian_model_ht.py:809-822)."""
    tanfovx = math.tan(fovx * 0.5)
    focal = W / (2.0 * tanfovx)
    tanfovy = H / (2.0 * focal)
    a = math.radians(yaw_deg)
    ca, sa = math.cos(a), math.sin(a)
    # world->camera rotation R and translation t such that the pivot stays on the optical axis
    R = torch.tensor([[ca, 0.0, -sa], [0.0, 1.0, 0.0], [sa, 0.0, ca]], dtype=torch.float64)
    pivot = torch.tensor([0.0, 0.0, pivot_z], dtype=torch.float64)
    t = pivot - R @ pivot
    w2c = torch.eye(4, dtype=torch.float64)
    w2c[:3, :3] = R
    w2c[:3, 3] = t
    view = w2c.t().contiguous().float()  # stored transposed (row-vector convention)
    proj = torch.zeros(4, 4, dtype=torch.float64)
    proj[0, 0] = 1.0 / tanfovx
    proj[1, 1] = 1.0 / tanfovy
    proj[3, 2] = 1.0
    proj[2, 2] = zfar / (zfar - znear)
    proj[2, 3] = -(zfar * znear) / (zfar - znear)
    proj_t = proj.t().contiguous().float()
    full = (view.unsqueeze(0).bmm(proj_t.unsqueeze(0))).squeeze(0).contiguous()
    campos = view.inverse()[3, :3].contiguous()
    return dict(W=W, H=H, tanfovx=tanfovx, tanfovy=tanfovy, viewmatrix=view.to(device), projmatrix=full.to(device),
                campos=campos.to(device), focal=focal)


def make_scene(P, W, H, s_max=8.0, seed=0, kind="default", fovx=FOVX_FRANCIS, device="cpu"):
    """Gaussian parameters as the HT-3DGS model hands them to the plugin (after activation):
    means3D[P,3], scales[P,3] (=exp(_scaling)), rotations[P,4] (normalised), opacities[P,1]
    (=sigmoid(_opacity)), shs[P,16,3]."""
    g = torch.Generator().manual_seed(int(seed))
    tanfovx = math.tan(fovx * 0.5)
    focal = W / (2.0 * tanfovx)
    tanfovy = H / (2.0 * focal)

    def U(n, lo, hi):
        return lo + (hi - lo) * torch.rand(n, generator=g)

    z = U(P, 1.0, 8.0)
    near = torch.rand(P, generator=g) < 0.02  # exercise the near-plane cull
    z = torch.where(near, U(P, -1.0, 0.2), z)
    x = z * tanfovx * U(P, -1.15, 1.15)  # ~13% off screen: exercises the rect clamp
    y = z * tanfovy * U(P, -1.15, 1.15)
    means3D = torch.stack([x, y, z], 1)
    if kind == "stress":
        s_px = U(P, 0.3, s_max)
        opacities = U(P, 0.02, 0.15).unsqueeze(1)
    else:
        s_px = torch.exp(U(P, math.log(0.4), math.log(s_max)))
        opacities = torch.sigmoid(2.0 * torch.randn(P, generator=g)).unsqueeze(1)
    base = s_px * z.abs().clamp_min(0.2) / focal
    aniso = torch.stack([torch.ones(P), U(P, 0.3, 1.0), U(P, 0.3, 1.0)], 1)
    scales = base.unsqueeze(1) * aniso
    q = torch.randn(P, 4, generator=g)
    rotations = q / q.norm(dim=1, keepdim=True)
    shs = torch.empty(P, 16, 3)
    shs[:, 0] = torch.randn(P, 3, generator=g)
    shs[:, 1:] = 0.3 * torch.randn(P, 15, 3, generator=g)
    out = dict(means3D=means3D, scales=scales, rotations=rotations, opacities=opacities, shs=shs)
    return {k: v.float().contiguous().to(device) for k, v in out.items()}


def make_upstream(W, H, seed=0, device="cpu"):
    """Fixed upstream gradients for the three outputs (all three backward paths live)."""
    g = torch.Generator().manual_seed(1000 + int(seed))
    n = float(W * H)
    out = dict(dL_dcolor=torch.randn(3, H, W, generator=g) / (3.0 * n), dL_ddepth=torch.randn(1, H, W, generator=g) / n,
               dL_dalpha=torch.randn(1, H, W, generator=g) / n)
    return {k: v.float().contiguous().to(device) for k, v in out.items()}


def make_config(name, device="cpu", **override):
    cfg = dict(CONFIGS[name])
    cfg.update(override)
    scene = make_scene(cfg["P"], cfg["W"], cfg["H"], cfg["s_max"], cfg["seed"], cfg["kind"], device=device)
    cam = make_camera(cfg["W"], cfg["H"], device=device)
    up = make_upstream(cfg["W"], cfg["H"], cfg["seed"], device=device)
    return cfg, scene, cam, up
