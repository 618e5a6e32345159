"""Where does the end-to-end step spend its extra time over the device-timed step?  Development aid."""
import importlib, os, sys, time
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
importlib.import_module("3dgs_hierarchical_training_b200").install()
import htgs_synthetic as syn
import diff_gaussian_rasterization as dgr

dev = torch.device("cuda", 0)
cfg, scene, cam_h, up = syn.make_config("config3", device="cpu")
P, W, H = cfg["P"], cfg["W"], cfg["H"]
params = {k: v.to(dev).requires_grad_(True) for k, v in scene.items()}
bg = torch.zeros(3, device=dev)
up_d = {k: v.to(dev) for k, v in up.items()}
cam_d = {k: (v.to(dev) if hasattr(v, "to") else v) for k, v in cam_h.items()}
host = {k: v.pin_memory() for k, v in up.items()}
host_cam = {k: cam_h[k].pin_memory() for k in ("viewmatrix", "projmatrix", "campos")}
result = torch.zeros(1, dtype=torch.float64).pin_memory()
copy_stream = torch.cuda.Stream(device=dev)
copied = torch.cuda.Event()


def settings(cam):
    return dgr.GaussianRasterizationSettings(image_height=H, image_width=W, tanfovx=cam["tanfovx"], tanfovy=cam["tanfovy"], bg=bg,
                                             scale_modifier=1.0, viewmatrix=cam["viewmatrix"], projmatrix=cam["projmatrix"],
                                             sh_degree=3, campos=cam["campos"], prefiltered=False, debug=False)


def step(cam, upg, hook=None):
    for p in params.values():
        p.grad = None
    m2 = torch.zeros_like(params["means3D"], requires_grad=True)
    color, radii, depth, alpha = dgr.GaussianRasterizer(settings(cam))(means3D=params["means3D"], means2D=m2, opacities=params["opacities"],
                                                                       shs=params["shs"], scales=params["scales"], rotations=params["rotations"])
    if hook:
        hook()
    torch.autograd.backward((color, depth, alpha), (upg["dL_dcolor"], upg["dL_ddepth"], upg["dL_dalpha"]))
    return color


def variant(h2d, readback, overlap=True):
    def f():
        main = torch.cuda.current_stream(dev)
        if h2d:
            if overlap:
                copy_stream.wait_stream(main)
                with torch.cuda.stream(copy_stream):
                    upg = {k: v.to(dev, non_blocking=True) for k, v in host.items()}
                    copied.record(copy_stream)
                hook = lambda: main.wait_event(copied)
            else:
                upg = {k: v.to(dev, non_blocking=True) for k, v in host.items()}
                hook = None
            cam = dict(cam_h)
            for k, v in host_cam.items():
                cam[k] = v.to(dev, non_blocking=True)
        else:
            upg, cam, hook = up_d, cam_d, None
        color = step(cam, upg, hook)
        if readback == "sum":
            s = (color.detach().sum(dtype=torch.float64) + params["opacities"].grad.sum(dtype=torch.float64)).reshape(1)
            result.copy_(s, non_blocking=False)
        elif readback == "sync":
            torch.cuda.synchronize()
    return f


def wall(f, n=100):
    for _ in range(5):
        f()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(n):
        f()
    torch.cuda.synchronize()
    return 1e3 * (time.perf_counter() - t0) / n


def cpu_only(n=100):
    f = variant(False, None)
    for _ in range(5):
        f()
    torch.cuda.synchronize()
    ts = []
    for _ in range(n):
        t0 = time.perf_counter()
        f()
        ts.append(time.perf_counter() - t0)
        torch.cuda.synchronize()
    return 1e3 * sorted(ts)[n // 2]


def variant_bg_copy(nbytes_scale):
    """device-resident inputs, but an unrelated H2D of the same size runs on the side stream every step (never waited for)"""
    base = variant(False, "sum")
    junk = [torch.empty_like(v, device=dev) for v in host.values()]
    def f():
        with torch.cuda.stream(copy_stream):
            for j, v in zip(junk, host.values()):
                n = int(v.numel() * nbytes_scale)
                j.view(-1)[:n].copy_(v.view(-1)[:n], non_blocking=True)
        base()
    return f


print("device-resident + unrelated 41 MB H2D  %.3f ms" % wall(variant_bg_copy(1.0)))
print("device-resident + unrelated tiny H2D   %.3f ms" % wall(variant_bg_copy(0.0001)))
print("device-resident, no per-step sync      %.3f ms" % wall(variant(False, None)))
print("device-resident, sync per step         %.3f ms" % wall(variant(False, "sync")))
print("device-resident, sum+readback per step %.3f ms" % wall(variant(False, "sum")))
print("H2D overlapped,  sum+readback per step %.3f ms" % wall(variant(True, "sum")))
print("H2D serial,      sum+readback per step %.3f ms" % wall(variant(True, "sum", overlap=False)))
print("host time to enqueue one step (median) %.3f ms" % cpu_only())
