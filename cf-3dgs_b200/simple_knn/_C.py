"""`distCUDA2(points)` with the reference's signature and result (submodules/simple-knn/spatial.cu:15-26, ext.cpp:15-17),
computed by libgsr_b200.so (csrc/gsr_knn.cu) through the C ABI (`gsr_knn_mean_dist2`, include/gsr.h).

    from simple_knn._C import distCUDA2          # scene/gaussian_model_cf.py:20, scene/gaussian_model.py:20
    dist2 = torch.clamp_min(distCUDA2(points), 1e-7)

There is no CPU path: a CPU tensor raises, like the reference's kernel launch on a host pointer would fail.
"""
from __future__ import annotations

import torch

from diff_gaussian_rasterization._cabi import check, lib


def distCUDA2(points: torch.Tensor) -> torch.Tensor:
    if not points.is_cuda:
        raise RuntimeError("distCUDA2: points must be a CUDA tensor (this build has no CPU path)")
    if points.dtype != torch.float32:
        # the reference calls points.data<float>(), which throws for any other dtype
        raise RuntimeError(f"distCUDA2: expected a float32 tensor, got {points.dtype}")
    if points.dim() != 2 or points.shape[1] != 3:
        raise RuntimeError(f"distCUDA2: expected points of shape (P, 3), got {tuple(points.shape)}")
    pts = points.contiguous()
    P = pts.shape[0]
    means = torch.zeros(P, dtype=torch.float32, device=pts.device)   # spatial.cu:21: torch::full({P}, 0.0)
    if P == 0:
        return means
    nbytes = lib.gsr_knn_scratch_bytes(P)
    scratch = torch.empty(nbytes, dtype=torch.uint8, device=pts.device)
    with torch.cuda.device(pts.device):
        check(lib.gsr_knn_mean_dist2(P, pts.data_ptr(), means.data_ptr(), scratch.data_ptr(), nbytes,
                                     torch.cuda.current_stream().cuda_stream))
    return means
