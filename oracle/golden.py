"""Python face of the CPU oracle (oracle/htgs_oracle.c -> liborc.so).

TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and
bench.py's cpu_baseline / --impl reference legs; never by the product.

Works on numpy arrays (fp32, C-contiguous).  `forward` returns every
intermediate the reference keeps in its Geometry/Binning/Image states so parity
tests can compare stage by stage; `backward` returns the eight gradients of
rasterize_points.cu:158-167 plus the internal dL_dconic / dL_ddepths.
"""
import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "liborc.so")
_lib = None

_f = ctypes.c_float
_i = ctypes.c_int
_p = ctypes.c_void_p
_i64 = ctypes.c_int64


def build(force=False):
    src = os.path.join(_HERE, "htgs_oracle.c")
    if force or not os.path.exists(_LIB_PATH) or os.path.getmtime(src) > os.path.getmtime(_LIB_PATH):
        subprocess.check_call(["make", "-C", _HERE, "liborc.so"], stdout=subprocess.DEVNULL)
    return _LIB_PATH


def lib():
    global _lib
    if _lib is None:
        build()
        l = ctypes.CDLL(_LIB_PATH)
        l.orc_preprocess.restype = _i64
        l.orc_preprocess.argtypes = [_i, _i, _i, _p, _p, _f, _p, _p, _p, _p, _p, _p, _p, _p, _i, _i, _f, _f,
                                     _p, _p, _p, _p, _p, _p, _p, _p]
        l.orc_bin.restype = None
        l.orc_bin.argtypes = [_i, _i, _i, _p, _p, _p, _p, _p, _i64, _p, _p, _p, _p, _p]
        l.orc_render.restype = None
        l.orc_render.argtypes = [_i, _i, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p]
        l.orc_render_backward.restype = None
        l.orc_render_backward.argtypes = [_i, _i, _i] + [_p] * 17
        l.orc_preprocess_backward.restype = None
        l.orc_preprocess_backward.argtypes = [_i, _i, _i, _p, _p, _p, _p, _p, _p, _f, _p, _p, _p, _p, _i, _i, _f, _f,
                                              _p, _p, _p, _p, _p, _p, _p, _p, _p]
        l.orc_mark_visible.restype = None
        l.orc_mark_visible.argtypes = [_i, _p, _p, _p, _p]
        l.orc_dist2.restype = None
        l.orc_dist2.argtypes = [_i, _p, _p]
        l.orc_higher_msb.restype = ctypes.c_uint32
        l.orc_higher_msb.argtypes = [ctypes.c_uint32]
        l.orc_num_threads.restype = _i
        l.orc_set_num_threads.argtypes = [_i]
        _lib = l
    return _lib


def num_threads():
    return int(lib().orc_num_threads())


def set_num_threads(n):
    lib().orc_set_num_threads(int(n))


def _a(x, dtype=np.float32):
    """numpy view of a tensor/array as C-contiguous `dtype`, or None for absent options."""
    if x is None:
        return None
    if hasattr(x, "detach"):
        x = x.detach().cpu().numpy()
    x = np.ascontiguousarray(x, dtype=dtype)
    return x if x.size else None


def _ptr(x):
    return None if x is None else x.ctypes.data_as(_p)


def forward(means3D, opacities, cam, bg, sh_degree=3, shs=None, colors_precomp=None, scales=None, rotations=None,
            cov3D_precomp=None, scale_modifier=1.0):
    """Oracle of Rasterizer::forward (rasterizer_impl.cu:197-339). `cam` is the dict of
    htgs_synthetic.make_camera (numpy or torch values)."""
    l = lib()
    means3D, opacities, shs, colors_precomp = _a(means3D), _a(opacities), _a(shs), _a(colors_precomp)
    scales, rotations, cov3D_precomp = _a(scales), _a(rotations), _a(cov3D_precomp)
    view, proj, campos, bg = _a(cam["viewmatrix"]), _a(cam["projmatrix"]), _a(cam["campos"]), _a(bg)
    W, H = int(cam["W"]), int(cam["H"])
    P = 0 if means3D is None else means3D.shape[0]
    M = 0 if shs is None else shs.shape[1]
    tiles = ((W + 15) // 16) * ((H + 15) // 16)
    st = dict(P=P, M=M, D=sh_degree, W=W, H=H, scale_modifier=float(scale_modifier),
              tanfovx=float(cam["tanfovx"]), tanfovy=float(cam["tanfovy"]))
    st["radii"] = np.zeros(P, np.int32)
    st["means2D"] = np.zeros((P, 2), np.float32)
    st["depths"] = np.zeros(P, np.float32)
    st["cov3D"] = np.zeros((P, 6), np.float32)
    st["rgb"] = np.zeros((P, 3), np.float32)
    st["conic_opacity"] = np.zeros((P, 4), np.float32)
    st["tiles_touched"] = np.zeros(P, np.uint32)
    st["clamped"] = np.zeros((P, 3), np.uint8)
    st["point_offsets"] = np.zeros(P, np.uint32)
    st["ranges"] = np.zeros((tiles, 2), np.uint32)
    st["n_contrib"] = np.zeros((H, W), np.uint32)
    st["color"] = np.zeros((3, H, W), np.float32)
    st["depth"] = np.zeros((1, H, W), np.float32)
    st["alpha"] = np.zeros((1, H, W), np.float32)
    st["counters"] = np.zeros(2, np.int64)
    R = 0
    if P > 0:
        R = int(l.orc_preprocess(P, sh_degree, M, _ptr(means3D), _ptr(scales), float(scale_modifier), _ptr(rotations),
                                 _ptr(opacities), _ptr(shs), _ptr(cov3D_precomp), _ptr(colors_precomp), _ptr(view),
                                 _ptr(proj), _ptr(campos), W, H, st["tanfovx"], st["tanfovy"], _ptr(st["radii"]),
                                 _ptr(st["means2D"]), _ptr(st["depths"]), _ptr(st["cov3D"]), _ptr(st["rgb"]),
                                 _ptr(st["conic_opacity"]), _ptr(st["tiles_touched"]), _ptr(st["clamped"])))
    st["R"] = R
    st["keys_unsorted"] = np.zeros(R, np.uint64)
    st["values_unsorted"] = np.zeros(R, np.uint32)
    st["keys"] = np.zeros(R, np.uint64)
    st["point_list"] = np.zeros(R, np.uint32)
    if P > 0:
        l.orc_bin(P, W, H, _ptr(st["radii"]), _ptr(st["means2D"]), _ptr(st["depths"]), _ptr(st["tiles_touched"]),
                  _ptr(st["point_offsets"]), R, _ptr(st["keys_unsorted"]) if R else None,
                  _ptr(st["values_unsorted"]) if R else None, _ptr(st["keys"]) if R else None,
                  _ptr(st["point_list"]) if R else None, _ptr(st["ranges"]))
        colors = colors_precomp if colors_precomp is not None else st["rgb"]
        l.orc_render(W, H, _ptr(st["ranges"]), _ptr(st["point_list"]) if R else None, _ptr(st["means2D"]), _ptr(colors),
                     _ptr(st["depths"]), _ptr(st["conic_opacity"]), _ptr(bg), _ptr(st["color"]), _ptr(st["depth"]),
                     _ptr(st["alpha"]), _ptr(st["n_contrib"]), _ptr(st["counters"]))
    # P == 0: images stay zero, background not applied (rasterize_points.cu:83)
    st["_inputs"] = dict(means3D=means3D, opacities=opacities, shs=shs, colors_precomp=colors_precomp, scales=scales,
                         rotations=rotations, cov3D_precomp=cov3D_precomp, view=view, proj=proj, campos=campos, bg=bg)
    return st


def backward(st, dL_dcolor, dL_ddepth, dL_dalpha):
    """Oracle of Rasterizer::backward (rasterizer_impl.cu:343-447) on the state of `forward`."""
    l = lib()
    inp = st["_inputs"]
    P, M, W, H = st["P"], st["M"], st["W"], st["H"]
    dL_dcolor, dL_ddepth, dL_dalpha = _a(dL_dcolor), _a(dL_ddepth), _a(dL_dalpha)
    g = dict(dL_dmeans2D=np.zeros((P, 3), np.float32), dL_dconic=np.zeros((P, 4), np.float32),
             dL_dopacity=np.zeros((P, 1), np.float32), dL_dcolors=np.zeros((P, 3), np.float32),
             dL_ddepths=np.zeros((P, 1), np.float32), dL_dmeans3D=np.zeros((P, 3), np.float32),
             dL_dcov3D=np.zeros((P, 6), np.float32), dL_dsh=np.zeros((P, M, 3), np.float32),
             dL_dscales=np.zeros((P, 3), np.float32), dL_drotations=np.zeros((P, 4), np.float32))
    if P == 0:
        return g
    colors = inp["colors_precomp"] if inp["colors_precomp"] is not None else st["rgb"]
    l.orc_render_backward(P, W, H, _ptr(st["ranges"]), _ptr(st["point_list"]) if st["R"] else None, _ptr(inp["bg"]),
                          _ptr(st["means2D"]), _ptr(st["conic_opacity"]), _ptr(colors), _ptr(st["depths"]),
                          _ptr(st["alpha"]), _ptr(st["n_contrib"]), _ptr(dL_dcolor), _ptr(dL_ddepth), _ptr(dL_dalpha),
                          _ptr(g["dL_dmeans2D"]), _ptr(g["dL_dconic"]), _ptr(g["dL_dopacity"]), _ptr(g["dL_dcolors"]),
                          _ptr(g["dL_ddepths"]))
    cov3D = inp["cov3D_precomp"] if inp["cov3D_precomp"] is not None else st["cov3D"]
    has_sh, has_sc = inp["shs"] is not None, inp["scales"] is not None
    l.orc_preprocess_backward(P, st["D"], M, _ptr(inp["means3D"]), _ptr(st["radii"]), _ptr(inp["shs"]),
                              _ptr(st["clamped"]), _ptr(inp["scales"]), _ptr(inp["rotations"]), st["scale_modifier"],
                              _ptr(cov3D), _ptr(inp["view"]), _ptr(inp["proj"]), _ptr(inp["campos"]), W, H,
                              st["tanfovx"], st["tanfovy"], _ptr(g["dL_dmeans2D"]), _ptr(g["dL_dconic"]),
                              _ptr(g["dL_dcolors"]), _ptr(g["dL_ddepths"]), _ptr(g["dL_dmeans3D"]), _ptr(g["dL_dcov3D"]),
                              _ptr(g["dL_dsh"]) if has_sh else None, _ptr(g["dL_dscales"]) if has_sc else None,
                              _ptr(g["dL_drotations"]) if has_sc else None)
    return g


def mark_visible(means3D, view, proj):
    means3D, view, proj = _a(means3D), _a(view), _a(proj)
    P = 0 if means3D is None else means3D.shape[0]
    out = np.zeros(P, np.uint8)
    if P:
        lib().orc_mark_visible(P, _ptr(means3D), _ptr(view), _ptr(proj), _ptr(out))
    return out.astype(bool)


def dist2(points):
    points = _a(points)
    P = 0 if points is None else points.shape[0]
    out = np.zeros(P, np.float32)
    if P:
        lib().orc_dist2(P, _ptr(points), _ptr(out))
    return out


def higher_msb(n):
    return int(lib().orc_higher_msb(int(n)))
