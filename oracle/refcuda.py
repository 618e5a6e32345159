"""ctypes door onto the UNMODIFIED reference CUDA code built into oracle/_ref/
(oracle/Makefile, oracle/ref_shim_dgr.cu, oracle/ref_shim_knn.cu).

TEST INFRASTRUCTURE ONLY (GPU parity tests, golden-vector generation, the A/B
timing arm of bench.py --impl refcuda).  Mirrors what rasterize_points.cu does
around CudaRasterizer::Rasterizer: allocate outputs with torch, pass raw pointers.
The reference launches on the legacy default stream, which is torch's default
current stream.
"""
import ctypes
import os

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
DGR_PATH = os.path.join(_HERE, "_ref", "libref_dgr.so")
KNN_PATH = os.path.join(_HERE, "_ref", "libref_knn.so")
CUB_PATH = os.path.join(_HERE, "_ref", "libref_cub.so")
_dgr = None
_knn = None
_p, _i, _f = ctypes.c_void_p, ctypes.c_int, ctypes.c_float


def available():
    return os.path.exists(DGR_PATH) and os.path.exists(KNN_PATH)


def dgr():
    global _dgr
    if _dgr is None:
        l = ctypes.CDLL(DGR_PATH)
        l.ref_forward.restype = _i
        l.ref_forward.argtypes = [_i, _i, _i, _p, _i, _i, _p, _p, _p, _p, _p, _f, _p, _p, _p, _p, _p, _f, _f, _i, _p, _p,
                                  _p, _p, _i]
        l.ref_backward.restype = None
        l.ref_backward.argtypes = [_i, _i, _i, _i, _p, _i, _i, _p, _p, _p, _p, _p, _f, _p, _p, _p, _p, _p, _f, _f] + [_p] * 14 + [_i]
        l.ref_mark_visible.restype = None
        l.ref_mark_visible.argtypes = [_i, _p, _p, _p, _p]
        l.ref_get.restype = ctypes.c_longlong
        l.ref_get.argtypes = [ctypes.c_char_p, _p, ctypes.c_size_t]
        l.ref_release.restype = None
        _dgr = l
    return _dgr


def knn():
    global _knn
    if _knn is None:
        l = ctypes.CDLL(KNN_PATH)
        l.ref_dist2.restype = None
        l.ref_dist2.argtypes = [_i, _p, _p]
        _knn = l
    return _knn


_cub = None


def cub_available():
    return os.path.exists(CUB_PATH)


def cub():
    """cub::DeviceRadixSort::SortPairs<u64,u32> as the reference calls it (oracle/ref_shim_cub.cu)."""
    global _cub
    if _cub is None:
        l = ctypes.CDLL(CUB_PATH)
        l.ref_cub_temp_bytes.restype = ctypes.c_size_t
        l.ref_cub_temp_bytes.argtypes = [ctypes.c_int64]
        l.ref_cub_sort_pairs.restype = _i
        l.ref_cub_sort_pairs.argtypes = [ctypes.c_int64, _i, _p, _p, _p, _p, _p, ctypes.c_size_t]
        _cub = l
    return _cub


def cub_sort_pairs(keys, values, end_bit, out=None, temp=None):
    """Sorted copies (keys_out, values_out) of int64 keys / int32 values on bits [0, end_bit); default stream."""
    n = keys.numel()
    ko, vo = out if out is not None else (torch.empty_like(keys), torch.empty_like(values))
    if temp is None:
        temp = torch.empty(max(1, cub().ref_cub_temp_bytes(n)), dtype=torch.uint8, device=keys.device)
    rc = cub().ref_cub_sort_pairs(n, end_bit, keys.data_ptr(), values.data_ptr(), ko.data_ptr(), vo.data_ptr(), temp.data_ptr(),
                                  temp.numel())
    if rc != 0:
        raise RuntimeError(f"cub SortPairs failed with CUDA error {rc}")
    return ko, vo


def _ptr(t):
    return None if t is None or t.numel() == 0 else t.data_ptr()


_STATE_DTYPES = dict(depths=torch.float32, clamped=torch.uint8, means2D=torch.float32, cov3D=torch.float32,
                     conic_opacity=torch.float32, rgb=torch.float32, tiles_touched=torch.int32, point_offsets=torch.int32,
                     point_list=torch.int32, point_list_keys=torch.int64, point_list_unsorted=torch.int32,
                     point_list_keys_unsorted=torch.int64, ranges=torch.int32, n_contrib=torch.int32)


def forward(means3D, opacities, cam, bg, sh_degree=3, shs=None, colors_precomp=None, scales=None, rotations=None,
            cov3D_precomp=None, scale_modifier=1.0, want_state=True):
    dev = means3D.device
    P, W, H = means3D.shape[0], int(cam["W"]), int(cam["H"])
    M = 0 if shs is None else shs.shape[1]
    f32 = dict(dtype=torch.float32, device=dev)
    out = dict(color=torch.zeros((3, H, W), **f32), depth=torch.zeros((1, H, W), **f32), alpha=torch.zeros((1, H, W), **f32),
               radii=torch.zeros((P,), dtype=torch.int32, device=dev), P=P, M=M, W=W, H=H)
    R = dgr().ref_forward(P, sh_degree, M, _ptr(bg), W, H, _ptr(means3D), _ptr(shs), _ptr(colors_precomp), _ptr(opacities),
                          _ptr(scales), float(scale_modifier), _ptr(rotations), _ptr(cov3D_precomp), _ptr(cam["viewmatrix"]),
                          _ptr(cam["projmatrix"]), _ptr(cam["campos"]), float(cam["tanfovx"]), float(cam["tanfovy"]), 0,
                          out["color"].data_ptr(), out["depth"].data_ptr(), out["alpha"].data_ptr(), out["radii"].data_ptr(), 0)
    out["R"] = int(R)
    if want_state:
        tiles = ((W + 15) // 16) * ((H + 15) // 16)
        sizes = dict(depths=P, clamped=3 * P, means2D=2 * P, cov3D=6 * P, conic_opacity=4 * P, rgb=3 * P, tiles_touched=P,
                     point_offsets=P, point_list=R, point_list_keys=R, point_list_unsorted=R, point_list_keys_unsorted=R,
                     ranges=2 * tiles, n_contrib=W * H)
        for name, n in sizes.items():
            t = torch.zeros((max(n, 1),), dtype=_STATE_DTYPES[name], device=dev)
            got = dgr().ref_get(name.encode(), t.data_ptr(), t.numel() * t.element_size())
            if got < 0:
                raise RuntimeError(f"ref_get({name}) failed")
            out[name] = t[:n]
    return out


def backward(fw, means3D, cam, bg, dL_dcolor, dL_ddepth, dL_dalpha, sh_degree=3, shs=None, colors_precomp=None,
             scales=None, rotations=None, cov3D_precomp=None, scale_modifier=1.0):
    """Must directly follow the `forward` whose state it uses (the shim keeps one state)."""
    dev = means3D.device
    P, M, W, H = fw["P"], fw["M"], fw["W"], fw["H"]
    z = lambda *s: torch.zeros(s, dtype=torch.float32, device=dev)
    g = dict(dL_dmeans2D=z(P, 3), dL_dconic=z(P, 2, 2), dL_dopacity=z(P, 1), dL_dcolors=z(P, 3), dL_ddepths=z(P, 1),
             dL_dmeans3D=z(P, 3), dL_dcov3D=z(P, 6), dL_dsh=z(P, M, 3), dL_dscales=z(P, 3), dL_drotations=z(P, 4))
    dgr().ref_backward(P, sh_degree, M, fw["R"], _ptr(bg), W, H, _ptr(means3D), _ptr(shs), _ptr(colors_precomp),
                       fw["alpha"].data_ptr(), _ptr(scales), float(scale_modifier), _ptr(rotations), _ptr(cov3D_precomp),
                       _ptr(cam["viewmatrix"]), _ptr(cam["projmatrix"]), _ptr(cam["campos"]), float(cam["tanfovx"]),
                       float(cam["tanfovy"]), fw["radii"].data_ptr(), _ptr(dL_dcolor), _ptr(dL_ddepth), _ptr(dL_dalpha),
                       g["dL_dmeans2D"].data_ptr(), g["dL_dconic"].data_ptr(), g["dL_dopacity"].data_ptr(),
                       g["dL_dcolors"].data_ptr(), g["dL_ddepths"].data_ptr(), g["dL_dmeans3D"].data_ptr(),
                       g["dL_dcov3D"].data_ptr(), _ptr(g["dL_dsh"]) or g["dL_dmeans3D"].data_ptr(),
                       g["dL_dscales"].data_ptr(), g["dL_drotations"].data_ptr(), 0)
    return g


def mark_visible(means3D, view, proj):
    P = means3D.shape[0]
    out = torch.zeros((P,), dtype=torch.bool, device=means3D.device)
    if P:
        dgr().ref_mark_visible(P, means3D.data_ptr(), view.data_ptr(), proj.data_ptr(), out.data_ptr())
    return out


def dist2(points):
    P = points.shape[0]
    out = torch.zeros((P,), dtype=torch.float32, device=points.device)
    if P:
        knn().ref_dist2(P, points.contiguous().data_ptr(), out.data_ptr())
    return out
