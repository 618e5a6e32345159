"""CPU-side checks of the boundary: the C-ABI library loads and exports every symbol include/htgs.h
declares, size/layout queries behave, and the plugin mirrors the reference's interface and error
behaviour (diff_gaussian_rasterization/__init__.py:160-223).  No compute is launched."""
import inspect
import os
import re

import pytest
import torch

import htgs_native as nat

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    txt = open(os.path.join(ROOT, "include", "htgs.h")).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    return sorted(set(re.findall(r"\b(htgs_[a-z0-9_]+)\s*\(", txt)))


def test_every_declared_symbol_is_exported_and_bound():
    lib = nat.lib()
    names = _declared()
    assert len(names) >= 18
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/htgs.h but not exported by libhtgs.so"
    assert set(names) == set(nat.EXPORTS), set(names) ^ set(nat.EXPORTS)
    assert lib.htgs_abi_version() == 1


def test_size_queries_monotone_and_aligned():
    lib = nat.lib()
    assert lib.htgs_geom_bytes(0) > 0
    prev = 0
    for P in (1, 255, 256, 257, 10_000, 1_000_000):
        b = lib.htgs_geom_bytes(P)
        assert b >= prev and b >= 49 * P
        prev = b
        lay = nat.geom_layout(P)
        assert all(v % 256 == 0 for v in lay.values()) and max(lay.values()) < b
    assert lib.htgs_binning_bytes(1_000_000, 1920, 1080) >= 24 * 1_000_000
    assert lib.htgs_image_bytes(1920, 1080) >= 4 * 1920 * 1080 + 8 * 120 * 68
    assert lib.htgs_backward_scratch_bytes(1000) >= 48 * 1000
    assert lib.htgs_sort_scratch_bytes(1 << 20) > 0
    assert lib.htgs_dist2_workspace_bytes(100_000) > 24 * 100_000


def test_null_pointer_calls_fail_with_message_not_crash():
    lib = nat.lib()
    import ctypes
    r = ctypes.c_int64(0)
    st = lib.htgs_forward_preprocess(10, 3, 16, None, None, 1.0, None, None, None, None, None, None, None, None, 64, 64,
                                     0.5, 0.5, 0, None, None, ctypes.byref(r), None)
    assert st != 0 and b"null" in lib.htgs_last_error()
    st = lib.htgs_forward_preprocess(0, 3, 16, None, None, 1.0, None, None, None, None, None, None, None, None, 64, 64,
                                     0.5, 0.5, 0, None, None, ctypes.byref(r), None)
    assert st == 0 and r.value == 0  # P == 0 is a no-op, as in rasterize_points.cu:83


def test_plugin_surface_matches_reference():
    import diff_gaussian_rasterization as d
    assert d.GaussianRasterizationSettings._fields == (
        "image_height", "image_width", "tanfovx", "tanfovy", "bg", "scale_modifier", "viewmatrix", "projmatrix",
        "sh_degree", "campos", "prefiltered", "debug")
    assert list(inspect.signature(d.GaussianRasterizer.forward).parameters) == [
        "self", "means3D", "means2D", "opacities", "shs", "colors_precomp", "scales", "rotations", "cov3D_precomp"]
    assert list(inspect.signature(d.rasterize_gaussians).parameters) == [
        "means3D", "means2D", "sh", "colors_precomp", "opacities", "scales", "rotations", "cov3Ds_precomp", "raster_settings"]
    assert issubclass(d._RasterizeGaussians, torch.autograd.Function)
    assert issubclass(d.GaussianRasterizer, torch.nn.Module)
    for fn in ("rasterize_gaussians", "rasterize_gaussians_backward", "mark_visible"):
        assert callable(getattr(d._C, fn))
    from simple_knn._C import distCUDA2
    assert callable(distCUDA2)
    t = d.cpu_deep_copy_tuple((torch.ones(2), 3, "x"))
    assert torch.equal(t[0], torch.ones(2)) and t[1:] == (3, "x")


def _rasterizer():
    import diff_gaussian_rasterization as d
    z = torch.zeros(3)
    rs = d.GaussianRasterizationSettings(16, 16, 1.0, 1.0, z, 1.0, torch.eye(4), torch.eye(4), 3, z, False, False)
    return d.GaussianRasterizer(rs)


def test_option_validation_raises_like_reference():
    r = _rasterizer()
    m = torch.zeros(4, 3)
    o = torch.ones(4, 1)
    with pytest.raises(Exception, match="excatly one of either SHs or precomputed colors"):
        r(m, m, o, scales=m, rotations=torch.zeros(4, 4))
    with pytest.raises(Exception, match="excatly one of either SHs or precomputed colors"):
        r(m, m, o, shs=torch.zeros(4, 16, 3), colors_precomp=m, scales=m, rotations=torch.zeros(4, 4))
    with pytest.raises(Exception, match="exactly one of either scale/rotation pair or precomputed 3D covariance"):
        r(m, m, o, colors_precomp=m)
    with pytest.raises(Exception, match="exactly one of either scale/rotation pair or precomputed 3D covariance"):
        r(m, m, o, colors_precomp=m, scales=m, rotations=torch.zeros(4, 4), cov3D_precomp=torch.zeros(4, 6))
    with pytest.raises(Exception, match="exactly one of either scale/rotation pair"):
        r(m, m, o, colors_precomp=m, scales=m)  # scales without rotations


def test_no_cpu_fallback():
    """CPU tensors must raise: the product has no CPU path (and must never route through oracle/)."""
    r = _rasterizer()
    m = torch.zeros(4, 3)
    with pytest.raises(RuntimeError, match="no CPU path"):
        r(m, m, torch.ones(4, 1), colors_precomp=m, scales=m, rotations=torch.zeros(4, 4))
    with pytest.raises(RuntimeError, match="means3D must have dimensions"):
        r(torch.zeros(4, 2), m, torch.ones(4, 1), colors_precomp=m, scales=m, rotations=torch.zeros(4, 4))
    from simple_knn._C import distCUDA2
    with pytest.raises(RuntimeError, match="no CPU path"):
        distCUDA2(torch.zeros(8, 3))
    src = open(os.path.join(ROOT, "3dgs_hierarchical_training_b200", "diff_gaussian_rasterization", "_C.py")).read()
    src += open(os.path.join(ROOT, "3dgs_hierarchical_training_b200", "htgs_native.py")).read()
    assert "oracle" not in src.replace("no CPU or PyTorch fallback", "")
