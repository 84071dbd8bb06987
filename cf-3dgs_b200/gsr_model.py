"""Fused Gaussian-model step: activations, their backward, Adam and densification statistics as CUDA kernels
(reference: scene/gaussian_model_cf.py:41-56,115-172,242-268,716-719).

The reference keeps six parameter tensors in a torch.optim.Adam with per-group learning rates and lets autograd run
sigmoid / exp / normalize / cat forward and backward around every rasterizer call.  FusedGaussianModel keeps the same
raw parameters (features as ONE (P,M,3) tensor, so no cat), produces the activated tensors the rasterizer wants with one
kernel, and applies the whole update -- chain rule through the activations + Adam + statistics -- with two more.
"""
from __future__ import annotations

import ctypes
from typing import Dict, Optional

import torch

from diff_gaussian_rasterization._cabi import check, lib

LR_ORDER = ("xyz", "f_dc", "f_rest", "opacity", "scaling", "rotation")


class FusedGaussianModel:
    def __init__(self, xyz, features, opacity_raw, log_scales, rot_raw, lrs: Dict[str, float], betas=(0.9, 0.999),
                 eps: float = 1e-15, device_step: bool = False):
        self.xyz = xyz.detach().contiguous().float().clone()
        # While the model is trained at SH degree 0, coefficient 0 lives in a dense (P, 1, 3) plane of its own (dc_mode,
        # see shs()); `features` is then brought up to date whenever it is read.
        self.features_dc, self.dc_mode, self._dc_storage = None, False, None
        self._features = features.detach().contiguous().float().clone()       # (P, M, 3), coefficient 0 = f_dc
        self.opacity_raw = opacity_raw.detach().contiguous().float().clone()  # (P, 1)
        self.log_scales = log_scales.detach().contiguous().float().clone()    # (P, 3)
        self.rot_raw = rot_raw.detach().contiguous().float().clone()          # (P, 4)
        dev = self.xyz.device
        self.P, self.M = self.xyz.shape[0], self.features.shape[1]
        self.lrs = dict(lrs)
        self.betas, self.eps, self.step_count = betas, eps, 0
        self.state = torch.zeros(lib.gsr_gaussian_adam_state_floats(self.P, self.M), dtype=torch.float32, device=dev)
        self.xyz_gradient_accum = torch.zeros(self.P, device=dev)
        self.denom = torch.zeros(self.P, device=dev)
        self.max_radii2D = torch.zeros(self.P, device=dev)
        self._act = None
        # device_step: the step number lives in device memory (gsr_gaussian_adam_step_dev), so step() can be captured in
        # a CUDA graph and replayed; step_count on the host is then only a call counter
        self.step_state = torch.zeros(2, dtype=torch.int64, device=dev) if device_step else None
        # SH bands that have ever received a gradient (as a coefficient count): the bands above it still have zero moments,
        # so Adam's update of them is exactly zero and the step skips them (at degree 0: 3 of 48 floats per Gaussian)
        self.sh_ever = 0
        # set by gsr_parallel.BatchedViewsStep when the parameters live in symmetric memory and the step is shared between
        # the replicas: the moments then have the padded layout and only that object may step the model
        self.shared_step_owner = None

    @property
    def features(self) -> torch.Tensor:
        """(P, M, 3) SH coefficients.  In dc_mode coefficient 0 is copied in from the dense plane first (a strided write of
        12 bytes per Gaussian: for evaluation, densification, checkpoints -- not for the training loop, which uses shs())."""
        if self.dc_mode:
            self._features[:, 0].copy_(self.features_dc[:, 0])
        return self._features

    @features.setter
    def features(self, t: torch.Tensor) -> None:
        self._features = t
        self.features_dc, self.dc_mode = None, False

    def shs(self, sh_degree: int) -> torch.Tensor:
        """The tensor to hand to the rasterizer as `shs` at this SH degree.  At degree 0 only coefficient 0 is read, and of
        a (P, 16, 3) tensor that is 12 bytes out of every 192 -- for the rasterizer, for the SH gradient it writes back and
        for Adam.  So at degree 0 (every CF-3DGS local model, the global one's first thousand iterations) the model hands
        out a dense (P, 1, 3) plane that holds coefficient 0 and is what step() updates (dc_mode); the first call with a
        higher degree copies it back into `features` and leaves the mode.  The reference keeps _features_dc and
        _features_rest apart for its optimiser groups and concatenates them for every render
        (scene/gaussian_model_cf.py:130-160)."""
        if int(sh_degree) == 0 and self.M > 1:
            if not self.dc_mode:
                src = self._features[:, :1]
                if self._dc_storage is not None:
                    self._dc_storage.copy_(src)
                    self.features_dc = self._dc_storage
                else:
                    self.features_dc = src.contiguous()
                self.dc_mode = True
            return self.features_dc
        self._leave_dc()
        return self._features

    def _leave_dc(self) -> None:
        if self.dc_mode:
            self._features[:, 0].copy_(self.features_dc[:, 0])
            self.features_dc, self.dc_mode = None, False

    def activated(self) -> Dict[str, torch.Tensor]:
        """opacities (P,1), scales (P,3), rotations (P,4) for the rasterizer (fresh tensors, no autograd history)."""
        dev = self.xyz.device
        opa = torch.empty((self.P, 1), device=dev)
        sc = torch.empty((self.P, 3), device=dev)
        rot = torch.empty((self.P, 4), device=dev)
        with torch.cuda.device(dev):
            check(lib.gsr_gaussian_activate(self.P, self.opacity_raw.data_ptr(), self.log_scales.data_ptr(),
                                            self.rot_raw.data_ptr(), opa.data_ptr(), sc.data_ptr(), rot.data_ptr(),
                                            torch.cuda.current_stream().cuda_stream))
        return dict(opacities=opa, scales=sc, rotations=rot)

    def step(self, d_xyz, d_features, d_opacities, d_scales, d_rotations, d_means2D: Optional[torch.Tensor] = None,
             radii: Optional[torch.Tensor] = None, sh_degree: Optional[int] = None) -> None:
        """One optimiser step from the gradients the rasterizer backward produced (w.r.t. the activated tensors).
        sh_degree (needs device_step=True): the active SH degree the gradient was produced at -- the coefficients of higher
        bands have gradient zero; as long as they never had another, their moments are zero and they are skipped."""
        if self.shared_step_owner is not None:
            raise RuntimeError("this model is stepped by a BatchedViewsStep (shared step): call that object")
        if self.dc_mode and not (self.step_state is not None and d_features is not None and d_features.shape[1] == 1):
            self._leave_dc()      # a gradient for the full tensor: coefficient 0 goes back where that gradient points
        if d_features is not None and d_features.shape[1] == 1 and self.M > 1:
            if not self.dc_mode:
                raise ValueError("a (P,1,3) SH gradient belongs to the tensor shs(0) hands out (device_step=True): render "
                                 "from model.shs(0)")
            sh_degree = 0
        if sh_degree is not None and self.step_state is not None and d_features is not None:
            grads = dict(means3D=d_xyz.contiguous(), sh=d_features.contiguous(), opacity=d_opacities.contiguous(),
                         scales=d_scales.contiguous(), rotations=d_rotations.contiguous())
            self.step_part(3, grads, d_features.shape[1], tick=True, d_means2D=d_means2D, radii=radii,
                           active=(int(sh_degree) + 1) ** 2)
            return
        self.sh_ever = self.M
        self.step_count += 1
        lr = (ctypes.c_float * 6)(*[float(self.lrs[k]) for k in LR_ORDER])
        stats = d_means2D is not None and radii is not None
        p = lambda t: None if t is None else t.contiguous().data_ptr()
        tail = (self.xyz.data_ptr(), self._features.data_ptr(), self.opacity_raw.data_ptr(), self.log_scales.data_ptr(),
                self.rot_raw.data_ptr(), p(d_xyz), p(d_features), p(d_opacities), p(d_scales), p(d_rotations),
                self.state.data_ptr(), p(d_means2D) if stats else None, p(radii) if stats else None,
                self.xyz_gradient_accum.data_ptr() if stats else None, self.denom.data_ptr() if stats else None,
                self.max_radii2D.data_ptr() if stats else None, torch.cuda.current_stream().cuda_stream)
        with torch.cuda.device(self.xyz.device):
            if self.step_state is not None:
                check(lib.gsr_gaussian_adam_step_dev(self.P, self.M, self.step_state.data_ptr(), float(self.betas[0]),
                                                     float(self.betas[1]), float(self.eps), lr, *tail))
            else:
                check(lib.gsr_gaussian_adam_step(self.P, self.M, self.step_count, float(self.betas[0]),
                                                 float(self.betas[1]), float(self.eps), lr, *tail))

    def step_part(self, what: int, grads: Dict[str, torch.Tensor], M_grad: int, row0: int = 0, rows: Optional[int] = None,
                  tick: bool = False, d_means2D: Optional[torch.Tensor] = None, radii: Optional[torch.Tensor] = None,
                  active: Optional[int] = None) -> None:
        """A part of one optimiser step (gsr_gaussian_adam_step_part): what & 1 = the per-Gaussian block, what & 2 = the
        SH rows [row0, row0 + rows).  grads: "means3D" (P,3), "sh" (P,M_grad,3), "opacity" (P,1), "scales" (P,3),
        "rotations" (P,4) -- typically views into one flat, all-reduced buffer.  Needs device_step=True."""
        if self.step_state is None:
            raise RuntimeError("step_part needs device_step=True (the step counter lives on the device)")
        if self.shared_step_owner is not None:
            raise RuntimeError("this model is stepped by a BatchedViewsStep (shared step): call that object")
        rows = self.P if rows is None else rows
        # dc_mode: the gradient is that of the dense coefficient-0 plane; the same call with M = 1 steps it (the moment
        # arrays of coefficient 0 sit in front of the other SH moments, where a one-coefficient model has its own)
        dc = self.dc_mode and int(M_grad) == 1 and (active is None or int(active) == 1)
        if self.dc_mode and not dc:
            self._leave_dc()
        M_call, feat = (1, self.features_dc) if dc else (self.M, self._features)
        # active: coefficients that can hold a non-zero gradient in this step (default: all M_grad of them)
        self.sh_ever = max(self.sh_ever, min(int(M_grad), self.M) if active is None else min(int(active), self.M))
        if tick:
            self.step_count += 1
        lr = (ctypes.c_float * 6)(*[float(self.lrs[k]) for k in LR_ORDER])
        stats = bool(what & 1) and d_means2D is not None and radii is not None
        with torch.cuda.device(self.xyz.device):
            check(lib.gsr_gaussian_adam_step_part(
                self.P, M_call, int(M_grad), min(int(self.sh_ever), M_call), int(what), int(row0), int(rows), int(bool(tick)),
                self.step_state.data_ptr(),
                float(self.betas[0]), float(self.betas[1]), float(self.eps), lr,
                self.xyz.data_ptr(), feat.data_ptr(), self.opacity_raw.data_ptr(), self.log_scales.data_ptr(),
                self.rot_raw.data_ptr(), grads["means3D"].data_ptr(), grads["sh"].data_ptr(), grads["opacity"].data_ptr(),
                grads["scales"].data_ptr(), grads["rotations"].data_ptr(), self.state.data_ptr(),
                d_means2D.data_ptr() if stats else None, radii.data_ptr() if stats else None,
                self.xyz_gradient_accum.data_ptr() if stats else None, self.denom.data_ptr() if stats else None,
                self.max_radii2D.data_ptr() if stats else None, torch.cuda.current_stream().cuda_stream))
