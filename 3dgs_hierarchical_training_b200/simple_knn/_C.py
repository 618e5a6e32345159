"""Stand-in for the reference's pybind11 module `simple_knn._C`
(submodules/simple-knn/ext.cpp:15-17, spatial.cu:15-26)."""
import ctypes

import torch

import htgs_native as nat


def distCUDA2(points):
    """float32[P]: mean of the squared distances from each point to its 3 nearest other points.
    Same device/dtype contract as the reference (CUDA float tensor [P,3] in, [P] out)."""
    if not points.is_cuda:
        raise RuntimeError("distCUDA2: points must be a CUDA tensor (no CPU path)")
    if points.dtype != torch.float32:
        raise RuntimeError(f"expected scalar type Float but found {points.dtype}")
    dev = points.device
    P = int(points.size(0))
    means = torch.zeros((P,), dtype=torch.float32, device=dev)
    if P == 0:
        return means
    pts = points.contiguous()
    lib = nat.lib()
    with torch.cuda.device(dev):
        nbytes = lib.htgs_dist2_workspace_bytes(P)
        ws = torch.empty(nbytes, dtype=torch.uint8, device=dev)
        st = ctypes.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
        nat.check(lib.htgs_dist2(P, pts.data_ptr(), means.data_ptr(), ws.data_ptr(), nbytes, st), "htgs_dist2")
    return means
