"""`simple_knn._C.distCUDA2(points[P,3] float32 cuda) -> float32[P]`: mean squared distance to the three
nearest other points (reference submodules/simple-knn/spatial.cu:15-26, simple_knn.cu:186-221), computed by
libgsr_b200.so's gsr_knn_dist2."""
from __future__ import annotations

import ctypes as C

import torch

from .. import _lib


def distCUDA2(points: torch.Tensor) -> torch.Tensor:
    if not points.is_cuda:
        raise _lib.GsrError("distCUDA2 needs a CUDA tensor: there is no CPU path")
    lib = _lib.lib()
    pts = points.detach()
    if pts.dtype != torch.float32:
        pts = pts.float()
    pts = pts.contiguous()
    P = pts.size(0)
    out = torch.zeros((P,), dtype=torch.float32, device=pts.device)      # spatial.cu:19-20
    if P == 0:
        return out
    nbytes = lib.gsr_knn_workspace_bytes(P)
    tmp = torch.empty(nbytes, dtype=torch.uint8, device=pts.device)
    with torch.cuda.device(pts.device):
        rc = lib.gsr_knn_dist2(P, pts.data_ptr(), out.data_ptr(), tmp.data_ptr(), nbytes,
                               C.c_void_p(torch.cuda.current_stream(pts.device).cuda_stream))
    _lib.check(rc, "gsr_knn_dist2")
    return out
