"""``distCUDA2(points[P,3]) -> f32[P]``: mean squared distance to the 3 nearest neighbours.

Mirrors the pybind11 module of the reference (submodules/simple-knn/ext.cpp:14-16, spatial.cu:15-26) on top of
``vcr_dist2_knn3`` (include/vcr_gaus_b200.h).  CUDA only; no CPU fallback.
"""
import ctypes

import torch

import vcr_native as _n


def distCUDA2(points: torch.Tensor) -> torch.Tensor:
    if not points.is_cuda:
        raise RuntimeError("distCUDA2 (B200 build) needs a CUDA tensor; there is no CPU fallback")
    P = points.shape[0]
    pts = points.contiguous().to(torch.float32)
    means = torch.zeros((P,), dtype=torch.float32, device=points.device)
    if P == 0:
        return means
    scratch = torch.empty(int(_n.vcr_knn_scratch_bytes(P)), dtype=torch.uint8, device=points.device)
    _n.check(_n.vcr_dist2_knn3(P, pts.data_ptr(), means.data_ptr(), scratch.data_ptr(),
                               ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)), "distCUDA2")
    return means
