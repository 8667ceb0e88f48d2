"""Host-side mirror of `simple_knn._C.distCUDA2` (submodules/simple-knn/ext.cpp:15-17,
spatial.cu:15-26) on top of gsm_knn_dist2."""
import ctypes

import torch

from . import _lib


def distCUDA2(points):
    """Mean squared distance of each point to its 3 nearest other points: [P,3] f32 cuda -> [P] f32."""
    lib = _lib.load()
    if not points.is_cuda:
        raise RuntimeError("points must be a CUDA tensor")
    if points.dtype != torch.float32:
        raise RuntimeError(f"points must be float32, got {points.dtype}")
    P = points.shape[0]
    dev = points.device
    pts = points.contiguous()
    means = torch.full((P,), 0.0, dtype=torch.float32, device=dev)
    if P == 0:
        return means
    with torch.cuda.device(dev):
        ws = torch.empty((lib.gsm_knn_workspace_bytes(P),), dtype=torch.uint8, device=dev)
        stream = ctypes.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
        _lib.check(lib.gsm_knn_dist2(P, ctypes.c_void_p(pts.data_ptr()), ctypes.c_void_p(means.data_ptr()),
                                     ctypes.c_void_p(ws.data_ptr()), ws.numel(), stream), "knn_dist2")
    return means
