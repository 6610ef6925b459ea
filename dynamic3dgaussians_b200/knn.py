"""Host side of the kNN kernels: drop-in for simple_knn._C.distCUDA2 (K/spatial.cu:15-27) plus a general k."""
from __future__ import annotations

import torch

from . import _lib


def knn(points: torch.Tensor, k: int):
    """(mean_dist2 [P], idx [P,k] int32, dist2 [P,k]) of the k nearest OTHER points, ascending distance."""
    if not torch.is_tensor(points) or not points.is_cuda:
        raise ValueError("points must be a CUDA tensor (there is no CPU path)")
    if points.dim() != 2 or points.shape[1] != 3:
        raise ValueError("points must have shape [P, 3]")
    pts = points.detach().to(torch.float32).contiguous()
    P = pts.shape[0]
    dev = pts.device
    mean = torch.zeros(P, dtype=torch.float32, device=dev)
    idx = torch.zeros((P, k), dtype=torch.int32, device=dev)
    d2 = torch.zeros((P, k), dtype=torch.float32, device=dev)
    if P == 0:
        return mean, idx, d2
    lib = _lib.load()
    nbytes = lib.gsr_knn_temp_bytes(P)
    tmp = torch.empty(max(nbytes, 16), dtype=torch.uint8, device=dev)
    with torch.cuda.device(dev):
        _lib.check(lib.gsr_knn(P, k, pts.data_ptr(), mean.data_ptr(), idx.data_ptr(), d2.data_ptr(), tmp.data_ptr(),
                               nbytes, torch.cuda.current_stream(dev).cuda_stream), "knn")
    return mean, idx, d2


def distCUDA2(points: torch.Tensor):
    """As the vendored K/spatial.cu:15-27: (mean squared distance to the 3 nearest neighbours [P], their indices [P,3])."""
    mean, idx, _ = knn(points, 3)
    return mean, idx


def distCUDA2_dist(points: torch.Tensor) -> torch.Tensor:
    """Upstream single-tensor signature that REF/scene/gaussian_model.py:162 expects."""
    return knn(points, 3)[0]
