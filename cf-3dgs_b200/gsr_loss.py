"""Fused photometric loss for the CF-3DGS training step.

    loss = (1 - lambda_dssim) * L1(img, gt) + lambda_dssim * (1 - SSIM(img, gt))

with the reference's definitions (utils/loss_utils.py:14-62, trainer/losses.py:98-136, lambda_dssim = 0.2 from
arguments/__init__.py:113).  One forward and one backward CUDA kernel (csrc/gsr_loss.cu) replace the five grouped
conv2d calls, ~20 element-wise kernels and their autograd backward; `clamp=True` folds in the image.clamp(0,1) that
CF-3DGS applies to the render first (scene/gaussian_model_cf.py:885).
"""
from __future__ import annotations

import torch

from diff_gaussian_rasterization._cabi import check, lib


class _PhotoLoss(torch.autograd.Function):
    @staticmethod
    def forward(ctx, img, gt, lambda_dssim, clamp):
        if not (img.is_cuda and gt.is_cuda):
            raise RuntimeError("photometric_loss: CUDA tensors required (there is no CPU path)")
        img = img.contiguous().float()
        gt = gt.contiguous().float()
        if img.shape != gt.shape or img.dim() != 3:
            raise RuntimeError("photometric_loss: img and gt must both be (C,H,W)")
        C, H, W = img.shape
        with torch.cuda.device(img.device):
            scratch = torch.empty(lib.gsr_photo_loss_scratch_bytes(C, H, W), dtype=torch.uint8, device=img.device)
            out = torch.empty(3, dtype=torch.float32, device=img.device)
            check(lib.gsr_photo_loss_forward(C, H, W, img.data_ptr(), gt.data_ptr(), float(lambda_dssim), int(bool(clamp)),
                                             scratch.data_ptr(), out.data_ptr(), torch.cuda.current_stream().cuda_stream))
        ctx.save_for_backward(img, gt, scratch)
        ctx.cfg = (float(lambda_dssim), int(bool(clamp)))
        ctx.mark_non_differentiable(out)
        return out[0], out

    @staticmethod
    def backward(ctx, grad_loss, _grad_parts):
        img, gt, scratch = ctx.saved_tensors
        lam, clamp = ctx.cfg
        C, H, W = img.shape
        g = grad_loss.reshape(1).contiguous().float()
        dimg = torch.empty_like(img)
        with torch.cuda.device(img.device):
            check(lib.gsr_photo_loss_backward(C, H, W, img.data_ptr(), gt.data_ptr(), lam, clamp, scratch.data_ptr(),
                                              g.data_ptr(), dimg.data_ptr(), torch.cuda.current_stream().cuda_stream))
        return dimg, None, None, None


def photometric_loss(img: torch.Tensor, gt: torch.Tensor, lambda_dssim: float = 0.2, clamp: bool = False,
                     return_parts: bool = False):
    """Scalar loss tensor (0-dim, on the device).  With return_parts also a 3-vector (loss, l1, ssim)."""
    loss, parts = _PhotoLoss.apply(img, gt, lambda_dssim, clamp)
    return (loss, parts) if return_parts else loss
