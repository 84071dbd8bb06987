"""Seeded synthetic cameras and Gaussian scenes.

Shared by the tests, the smoke run and bench.py so every arm (CUDA path, CPU
oracle, reference CUDA build) sees byte-identical inputs.  Everything is
generated on the CPU in float32 with a fixed torch.Generator seed.

Camera conventions follow the reference (scene/cameras.py:57-74, the `is_co3d`
branch every CF-3DGS camera takes, trainer/trainer.py:497-500):
viewmatrix = W2C^T (so its flat memory is column-major W2C), projmatrix =
viewmatrix @ P^T with the intrinsics-based P, znear 0.01, zfar 100.
"""
from __future__ import annotations

import math
from typing import Dict, Optional

import numpy as np
import torch

ZNEAR, ZFAR = 0.01, 100.0
DEFAULT_FOVX_DEG = 79.0  # trainer/trainer.py:268-274 heuristic for unknown intrinsics


def _roty(deg: float) -> np.ndarray:
    a = math.radians(deg)
    c, s = math.cos(a), math.sin(a)
    return np.array([[c, 0, s], [0, 1, 0], [-s, 0, c]], dtype=np.float64)


def make_camera(W: int, H: int, fovx_deg: float = DEFAULT_FOVX_DEG, R: Optional[np.ndarray] = None,
                t: Optional[np.ndarray] = None) -> Dict[str, object]:
    """R, t: world-to-camera rotation (3x3) and translation (3)."""
    R = np.eye(3) if R is None else np.asarray(R, dtype=np.float64)
    t = np.zeros(3) if t is None else np.asarray(t, dtype=np.float64)
    fx = W / (2.0 * math.tan(math.radians(fovx_deg) / 2.0))
    fy = fx
    cx, cy = W / 2.0, H / 2.0
    w2c = np.eye(4)
    w2c[:3, :3] = R
    w2c[:3, 3] = t
    view = torch.tensor(np.float32(w2c)).transpose(0, 1).contiguous()
    far, near = ZFAR, ZNEAR
    proj = torch.tensor([[2 * fx / W, 0.0, -(W - 2 * cx) / W, 0.0],
                         [0.0, 2 * fy / H, -(H - 2 * cy) / H, 0.0],
                         [0.0, 0.0, far / (far - near), -(far * near) / (far - near)],
                         [0.0, 0.0, 1.0, 0.0]]).float().transpose(0, 1)
    full = (view.unsqueeze(0).bmm(proj.unsqueeze(0))).squeeze(0).contiguous()
    campos = view.inverse()[3, :3].contiguous()
    return dict(W=W, H=H, viewmatrix=view, projmatrix=full, campos=campos,
                tanfovx=W / (2.0 * fx), tanfovy=H / (2.0 * fy), fx=fx, fy=fy)


def second_view(W: int, H: int) -> Dict[str, object]:
    """Camera rotated 10 degrees about y and shifted (0.3,0,0): exercises the
    off-axis covariance clamp (forward.cu:82-87)."""
    return make_camera(W, H, R=_roty(10.0), t=np.array([0.3, 0.0, 0.0]))


def gvol_scale_for(W: int, H: int, P: int) -> float:
    """Base scale s0 for G-vol.  Chosen (by measuring with the CPU oracle) so the mean number
    of tiles per visible Gaussian is ~9 at 1920x1080; scaled with resolution so the footprint in
    tiles stays comparable."""
    return 0.0028 * (1920.0 / W)


def make_gvol(P: int, W: int, H: int, seed: int, s0: Optional[float] = None, sh_degree: int = 3,
              fovx_deg: float = DEFAULT_FOVX_DEG) -> Dict[str, torch.Tensor]:
    """G-vol: volumetric cloud inside (and slightly around) the frustum of the canonical camera."""
    g = torch.Generator().manual_seed(seed)
    tanx = math.tan(math.radians(fovx_deg) / 2.0)
    tany = tanx * H / W
    z = 0.5 + 9.5 * torch.rand(P, generator=g)
    x = (2 * torch.rand(P, generator=g) - 1) * 1.2 * tanx * z
    y = (2 * torch.rand(P, generator=g) - 1) * 1.2 * tany * z
    behind = torch.rand(P, generator=g) < 0.02
    z = torch.where(behind, -1.0 + 1.19 * torch.rand(P, generator=g), z)
    means = torch.stack([x, y, z], 1).float().contiguous()
    if s0 is None:
        s0 = gvol_scale_for(W, H, P)
    # footprint in pixels ~ fx*s/z: scale with depth so near and far splats cover similar areas
    scales = torch.exp(math.log(s0) + 0.5 * torch.randn(P, 3, generator=g)) * z.abs().clamp(min=0.5).unsqueeze(1)
    rot = torch.randn(P, 4, generator=g)
    rot = rot / rot.norm(dim=1, keepdim=True)
    opac = torch.sigmoid(1.5 * torch.randn(P, 1, generator=g))
    M = 16
    shs = torch.zeros(P, M, 3)
    shs[:, 0] = torch.randn(P, 3, generator=g)
    shs[:, 1:] = 0.1 * torch.randn(P, M - 1, 3, generator=g)
    return dict(means3D=means, scales=scales.float().contiguous(), rotations=rot.float().contiguous(),
                opacities=opac.float().contiguous(), shs=shs.float().contiguous(), sh_degree=sh_degree)


def _depth_and_rgb(W: int, H: int, g: torch.Generator):
    v, u = torch.meshgrid(torch.arange(H, dtype=torch.float32), torch.arange(W, dtype=torch.float32),
                          indexing="ij")
    un, vn = u / W, v / H
    ph = torch.rand(8, generator=g) * 6.28318
    z = 3.5 + 0.9 * torch.sin(2 * math.pi * 1.0 * un + ph[0]) + 0.7 * torch.sin(2 * math.pi * 1.5 * vn + ph[1]) \
        + 0.5 * torch.sin(2 * math.pi * (2.0 * un + 1.0 * vn) + ph[2]) + 0.3 * torch.sin(2 * math.pi * 3.0 * (un - vn) + ph[3])
    z = z - 0.8 * (un > 0.6).float()
    z = z.clamp(1.0, 6.0)
    rgb = torch.stack([0.5 + 0.5 * torch.sin(2 * math.pi * 4 * un + ph[4]),
                       0.5 + 0.5 * torch.sin(2 * math.pi * 3 * vn + ph[5]),
                       0.5 + 0.5 * torch.sin(2 * math.pi * (2 * un + 5 * vn) + ph[6])], 0)
    return z, rgb


def make_gpix(W: int, H: int, seed: int, target_P: int = 250_000,
              fovx_deg: float = DEFAULT_FOVX_DEG) -> Dict[str, torch.Tensor]:
    """G-pix ("Tanks-like"): one small isotropic Gaussian per k-th pixel of a smooth synthetic
    depth map, as CF-3DGS initialises local models (trainer/trainer.py:489-512,
    scene/gaussian_model_cf.py:178-212).  sh_degree 0 with M=16."""
    g = torch.Generator().manual_seed(seed)
    z, rgb = _depth_and_rgb(W, H, g)
    k = max(1, int(round(math.sqrt(W * H / float(target_P)))))
    fx = W / (2.0 * math.tan(math.radians(fovx_deg) / 2.0))
    vs = torch.arange(0, H, k)
    us = torch.arange(0, W, k)
    vv, uu = torch.meshgrid(vs, us, indexing="ij")
    zz = z[vv, uu]
    x = (uu.float() + 0.5 - W / 2.0) / fx * zz
    y = (vv.float() + 0.5 - H / 2.0) / fx * zz
    means = torch.stack([x, y, zz], -1).reshape(-1, 3).float().contiguous()
    P = means.shape[0]
    # mean distance to the 3 nearest neighbours on the pixel grid ~ k * z / fx
    nn = (k * zz / fx).reshape(-1, 1)
    scales = nn.repeat(1, 3).float().contiguous()
    rot = torch.zeros(P, 4)
    rot[:, 0] = 1.0
    opac = torch.full((P, 1), 0.1)
    col = rgb[:, vv, uu].reshape(3, -1).t()
    shs = torch.zeros(P, 16, 3)
    shs[:, 0] = (col - 0.5) / 0.28209479177387814
    return dict(means3D=means, scales=scales, rotations=rot.contiguous(), opacities=opac.contiguous(),
                shs=shs.contiguous(), sh_degree=0, target_rgb=rgb, target_depth=z)


def cotangents(W: int, H: int, seed: int, dense: bool = True) -> Dict[str, torch.Tensor]:
    g = torch.Generator().manual_seed(seed)
    gc = torch.randn(3, H, W, generator=g)
    if dense:
        gd = torch.randn(1, H, W, generator=g)
        ga = torch.randn(1, H, W, generator=g)
    else:
        gd = torch.zeros(1, H, W)
        ga = torch.zeros(1, H, W)
    return dict(grad_color=gc, grad_depth=gd, grad_alpha=ga)
