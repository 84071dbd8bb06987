"""Rigid-pose optimisation on top of the rasterizer, without lietorch.

CF-3DGS fits a relative camera pose by rendering `SE3.retr(P).act(xyz)` and letting lietorch's backward turn
dL/dmeans3D into a 6-vector (scene/gaussian_model_cf.py:123-133, 303-373).  Here the same quantity comes out of the
rasterizer's backward kernel directly (gsr_backward's dL_dpose): for the left perturbation  p'(xi) = Exp(xi) * T * x
at xi = 0,
        dL/dxi = [ sum_i g_i ,  sum_i p'_i x g_i ],      g_i = dL/dmeans3D_i,   xi = (translation, rotation),
which is lietorch's SE3 tangent ordering.  lietorch is not vendored with the reference and cannot be installed
offline, so this convention is checked against float64 autograd through a matrix-exponential SE3, not against
lietorch itself.
"""
from __future__ import annotations

import torch

import diff_gaussian_rasterization as D


def hat(w: torch.Tensor) -> torch.Tensor:
    z = torch.zeros((), dtype=w.dtype, device=w.device)
    return torch.stack([torch.stack([z, -w[2], w[1]]), torch.stack([w[2], z, -w[0]]), torch.stack([-w[1], w[0], z])])


def se3_exp(xi: torch.Tensor) -> torch.Tensor:
    """4x4 matrix exponential of the twist xi = (v, w)."""
    A = torch.zeros(4, 4, dtype=xi.dtype, device=xi.device)
    A = A.clone()
    A[:3, :3] = hat(xi[3:])
    A[:3, 3] = xi[:3]
    return torch.linalg.matrix_exp(A)


def se3_exp_closed(xi: torch.Tensor) -> torch.Tensor:
    """Closed-form SE(3) exponential (Rodrigues) made of element-wise tensor ops only, so it can sit inside a
    captured CUDA graph (no matrix_exp, no host branch).  Series coefficients take over below 1e-4 rad."""
    v, w = xi[:3], xi[3:]
    th2 = (w * w).sum()
    th = torch.sqrt(th2 + 1e-30)
    small = th2 < 1e-8
    A = torch.where(small, 1.0 - th2 / 6.0, torch.sin(th) / th)
    B = torch.where(small, 0.5 - th2 / 24.0, (1.0 - torch.cos(th)) / (th2 + 1e-30))
    C = torch.where(small, 1.0 / 6.0 - th2 / 120.0, (th - torch.sin(th)) / (th2 * th + 1e-30))
    K = hat(w)
    K2 = K @ K
    eye = torch.eye(3, dtype=xi.dtype, device=xi.device)
    R = eye + A * K + B * K2
    V = eye + B * K + C * K2
    top = torch.cat([R, (V @ v).unsqueeze(1)], 1)
    bottom = torch.cat([torch.zeros(1, 3, dtype=xi.dtype, device=xi.device),
                        torch.ones(1, 1, dtype=xi.dtype, device=xi.device)], 1)   # device fills only: graph-capturable
    return torch.cat([top, bottom], 0)


class _PoseRasterize(torch.autograd.Function):
    """Renders xyz moved by T: the per-Gaussian kernels apply p' = T[:3,:3] p + T[:3,3] themselves (gsr_*_rigid, include/gsr.h),
    so there is no transformed copy of the points and no PyTorch op on the P x 3 tensor in either direction.  The
    gradient w.r.t. the twist `xi` (evaluated at xi = 0) is the fused dL_dpose; the gradient w.r.t. xyz comes back from
    the kernel already rotated into the frame of the raw points."""

    @staticmethod
    def forward(ctx, xi, xyz, T, sh, colors_precomp, opacities, scales, rotations, cov3Ds_precomp, rs):
        rigid = T[:3, :4].to(torch.float32).contiguous()
        out = D._forward_impl(rs.bg, xyz, colors_precomp, opacities, scales, rotations, rs.scale_modifier,
                              cov3Ds_precomp, rs.viewmatrix, rs.projmatrix, rs.tanfovx, rs.tanfovy, rs.image_height,
                              rs.image_width, sh, rs.sh_degree, rs.campos, rs.prefiltered, rs.debug, rigid=rigid)
        R, color, depth, alpha, radii, geom, binning, img = out
        ctx.rs, ctx.R = rs, R
        ctx.set_materialize_grads(False)
        ctx.mark_non_differentiable(radii)
        ctx.save_for_backward(colors_precomp, xyz, scales, rotations, cov3Ds_precomp, radii, sh, geom, binning, img,
                              alpha, rigid)
        return color, radii, depth, alpha

    @staticmethod
    def backward(ctx, grad_color, grad_radii, grad_depth, grad_alpha):
        rs = ctx.rs
        colors_precomp, xyz, scales, rotations, cov3Ds_precomp, radii, sh, geom, binning, img, alpha, rigid = ctx.saved_tensors
        if grad_color is None:
            grad_color = torch.zeros((3, rs.image_height, rs.image_width), dtype=torch.float32, device=xyz.device)
        # inputs: xi, xyz, T, sh, colors_precomp, opacities, scales, rotations, cov3Ds_precomp, rs -- a pose-only fit needs
        # none of the per-Gaussian gradients, and then the kernel does not write them (236 bytes per Gaussian)
        need = ctx.needs_input_grad
        skip = {"means2D"} | ({"means3D"} if not need[1] else set()) | ({"sh"} if not need[3] else set()) | \
               ({"colors"} if not need[4] else set()) | ({"opacity"} if not need[5] else set()) | \
               ({"scales"} if not (need[6] or need[7]) else set()) | ({"cov3D"} if not need[8] else set())
        res = D._backward_impl(rs.bg, xyz, radii, colors_precomp, scales, rotations, rs.scale_modifier, cov3Ds_precomp,
                               rs.viewmatrix, rs.projmatrix, rs.tanfovx, rs.tanfovy, grad_color, grad_depth, grad_alpha,
                               sh, rs.sh_degree, rs.campos, geom, ctx.R, binning, img, alpha, rs.debug, want_pose_grad=True,
                               rigid=rigid, skip=skip)
        dmeans2D, dcolors, dopacity, dxyz, dcov3D, dsh, dscales, drots, dpose = res
        return dpose, dxyz, None, dsh, dcolors, dopacity, dscales, drots, dcov3D, None


def rasterize_with_pose(raster_settings, xi, xyz, T, opacities, shs=None, colors_precomp=None, scales=None,
                        rotations=None, cov3D_precomp=None):
    """Render the Gaussians `xyz` moved by the rigid transform T (4x4, float32, CUDA).  `xi` is a 6-vector of zeros
    with requires_grad=True; after backward xi.grad holds the pose gradient (translation first).  Update the pose with
    T <- se3_exp(-lr * xi.grad) @ T (or any optimiser on xi followed by re-centring)."""
    e = torch.Tensor([])
    return _PoseRasterize.apply(xi, xyz, T, e if shs is None else shs, e if colors_precomp is None else colors_precomp,
                                opacities, e if scales is None else scales, e if rotations is None else rotations,
                                e if cov3D_precomp is None else cov3D_precomp, raster_settings)


class PoseOptimizer:
    """Pose-only fit of a frozen set of Gaussians against target frames (the local pose fit of CF-3DGS,
    trainer/cf3dgs_trainer.py:262-285) with nothing but library kernels in the loop: render through the rigid entry
    points, fused L1 + SSIM loss and its gradient, backward that reduces the twist gradient only, and the Adam step +
    SE(3) retraction on the device (gsr_pose_adam_step).  No autograd graph, no PyTorch element-wise kernel, no host
    read-back: with the sync-free forward (diff_gaussian_rasterization.set_sync_free) a whole step can be captured in a
    CUDA graph and replayed.  `T` (4x4, float32, on the device) is the current pose and is updated in place."""

    def __init__(self, raster_settings, xyz, opacities, shs=None, colors_precomp=None, scales=None, rotations=None,
                 cov3D_precomp=None, lr: float = 1e-3, betas=(0.9, 0.999), eps: float = 1e-8, lambda_dssim: float = 0.2,
                 clamp: bool = True, T0=None):
        from diff_gaussian_rasterization._cabi import lib
        self._lib = lib
        self.rs = raster_settings
        dev = xyz.device
        e = torch.Tensor([])
        c = lambda t: e if t is None else t.detach().contiguous().float()
        self.xyz, self.opacities = c(xyz), c(opacities)
        self.shs, self.colors, self.scales, self.rotations, self.cov3D = c(shs), c(colors_precomp), c(scales), c(rotations), c(cov3D_precomp)
        if shs is not None and int(raster_settings.sh_degree) == 0 and self.shs.shape[1] > 1:
            # the Gaussians are frozen and degree 0 reads coefficient 0 only: keep that as a dense (P,1,3) tensor instead of
            # 12 bytes out of every row of the caller's (P,M,3)
            self.shs = self.shs[:, :1].contiguous()
        self.T = (torch.eye(4, device=dev) if T0 is None else T0.detach().to(dev).float()).contiguous().clone()
        self.state = torch.zeros(13, device=dev)
        self.lr, self.betas, self.eps, self.lam, self.clamp = float(lr), betas, float(eps), float(lambda_dssim), int(bool(clamp))
        H, W = raster_settings.image_height, raster_settings.image_width
        with torch.cuda.device(dev):
            self._scratch = torch.empty(lib.gsr_photo_loss_scratch_bytes(3, H, W), dtype=torch.uint8, device=dev)
        self._parts = torch.zeros(3, device=dev)       # loss, l1, ssim of the latest step
        self._dimg = torch.empty(3, H, W, device=dev)
        self._one = torch.ones(1, device=dev)
        self._skip = ("means2D", "means3D", "opacity", "sh", "colors", "scales", "cov3D")
        self.grad = None

    def reset(self, T0=None) -> None:
        self.state.zero_()
        if T0 is None:
            self.T.copy_(torch.eye(4, device=self.T.device))
        else:
            self.T.copy_(T0)

    def step(self, target: torch.Tensor) -> torch.Tensor:
        """One iteration against `target` (3,H,W).  Returns the loss as a 0-dim device tensor (a view that the next step
        overwrites)."""
        from diff_gaussian_rasterization._cabi import check
        rs, lib = self.rs, self._lib
        H, W = rs.image_height, rs.image_width
        R, color, depth, alpha, radii, geom, binning, img = D._forward_impl(
            rs.bg, self.xyz, self.colors, self.opacities, self.scales, self.rotations, rs.scale_modifier, self.cov3D,
            rs.viewmatrix, rs.projmatrix, rs.tanfovx, rs.tanfovy, H, W, self.shs, rs.sh_degree, rs.campos, False, rs.debug,
            rigid=self.T)
        with torch.cuda.device(self.xyz.device):
            s = torch.cuda.current_stream().cuda_stream
            check(lib.gsr_photo_loss_forward(3, H, W, color.data_ptr(), target.data_ptr(), self.lam, self.clamp,
                                             self._scratch.data_ptr(), self._parts.data_ptr(), s))
            check(lib.gsr_photo_loss_backward(3, H, W, color.data_ptr(), target.data_ptr(), self.lam, self.clamp,
                                              self._scratch.data_ptr(), self._one.data_ptr(), self._dimg.data_ptr(), s))
        res = D._backward_impl(rs.bg, self.xyz, radii, self.colors, self.scales, self.rotations, rs.scale_modifier, self.cov3D,
                               rs.viewmatrix, rs.projmatrix, rs.tanfovx, rs.tanfovy, self._dimg, None, None, self.shs,
                               rs.sh_degree, rs.campos, geom, R, binning, img, alpha, rs.debug, want_pose_grad=True,
                               rigid=self.T, skip=self._skip)
        self.grad = res[8]
        with torch.cuda.device(self.xyz.device):
            check(lib.gsr_pose_adam_step(self.grad.data_ptr(), self.state.data_ptr(), self.T.data_ptr(), self.lr,
                                         float(self.betas[0]), float(self.betas[1]), self.eps,
                                         torch.cuda.current_stream().cuda_stream))
        return self._parts[0]
