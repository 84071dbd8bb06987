"""GPU: gsr_pose_gradient_static -- the pose-gradient entry point that keeps its scratch inside the library.

Some users of this library cannot hand it a scratch buffer because they cannot allocate one: a C tracking loop
on a time-sliced thread must avoid cudaMalloc/cudaFree (they lock the driver and synchronise the device), and
a process whose GPU memory is already owned by a framework allocator or a fixed MPS/MIG partition may not get
a fresh cudaMalloc at all.  For them the block partials sit in a 384 KiB __device__ array compiled into
libgsr_b200.so, which limits P to GSR_POSE_STATIC_MAX_POINTS (4,194,304).

The test calls the new entry point through the C ABI, twice in a row on different problems (the second one
smaller, so stale partials from the first call would show up if they were read), and compares each result
with what the existing entry point gsr_backward writes to dL_dpose for the same gradients.  The file name
sorts last so a `-x` run has already exercised every other test when it gets here; it also passes on its own
in a fresh process (the first call on a device does the one-time address look-up).
"""
import ctypes

import numpy as np
import pytest
import torch

import gsr_synth as S
from helpers import cuda_settings

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
MAX_POINTS = 4194304


def _existing_pose_gradient(P, W, H, seed):
    """dL/dmeans3D and dL/dpose from the existing entry point (gsr_backward with a caller-provided scratch)."""
    import diff_gaussian_rasterization as D
    cam = S.second_view(W, H)
    sc = S.make_gvol(P, W, H, seed=seed, s0=0.02)
    rs = cuda_settings(D, cam, (0.1, 0.2, 0.3), 3, DEV)
    t = {k: sc[k].to(DEV) for k in ["means3D", "opacities", "shs", "scales", "rotations"]}
    e = torch.Tensor([])
    R, color, depth, alpha, radii, geom, binning, img = D._forward_impl(
        rs.bg, t["means3D"], e, t["opacities"], t["scales"], t["rotations"], 1.0, e, rs.viewmatrix, rs.projmatrix,
        rs.tanfovx, rs.tanfovy, H, W, t["shs"], 3, rs.campos, False, False)
    ct = S.cotangents(W, H, seed=seed + 1, dense=True)
    out = D._backward_impl(rs.bg, t["means3D"], radii, e, t["scales"], t["rotations"], 1.0, e, rs.viewmatrix,
                           rs.projmatrix, rs.tanfovx, rs.tanfovy, ct["grad_color"].to(DEV), ct["grad_depth"].to(DEV),
                           ct["grad_alpha"].to(DEV), t["shs"], 3, rs.campos, geom, R, binning, img, alpha, False,
                           want_pose_grad=True)
    dmeans3D, dpose = out[3], out[8]
    torch.cuda.synchronize()
    return t["means3D"], dmeans3D, dpose


def _static_pose_gradient(means3D, dmeans3D):
    from diff_gaussian_rasterization import _cabi
    out = torch.full((6,), float("nan"), device=DEV)
    status = _cabi.lib.gsr_pose_gradient_static(means3D.shape[0], means3D.data_ptr(), dmeans3D.data_ptr(), out.data_ptr(),
                                                torch.cuda.current_stream().cuda_stream)
    assert status == 0, _cabi.lib.gsr_last_error()
    torch.cuda.synchronize()
    return out


def test_static_scratch_pose_gradient_matches_gsr_backward_twice():
    cases = [(60000, 240, 136, 71), (7001, 120, 72, 72)]     # second call: fewer blocks than the first
    for P, W, H, seed in cases:
        means3D, dmeans3D, want = _existing_pose_gradient(P, W, H, seed)
        got = _static_pose_gradient(means3D, dmeans3D)
        assert bool(torch.isfinite(got).all())
        err = float((got - want).abs().max()) / (float(want.abs().max()) + 1e-30)
        assert err <= 1e-4, (P, err, got.tolist(), want.tolist())
        # and against a float64 evaluation of the definition
        p, g = means3D.double(), dmeans3D.double()
        ref = torch.cat([g.sum(0), torch.cross(p, g, dim=1).sum(0)])
        err64 = float((got.double() - ref).abs().max()) / (float(ref.abs().max()) + 1e-30)
        assert err64 <= 1e-4, (P, err64)


def test_static_scratch_after_the_rigid_entry_points_needs_the_moved_points():
    """gsr_backward_rigid hands back dL/dxyz in the frame of the raw points (R^T g').  The static entry point wants
    the points the rasterizer saw, p' = R p + t, and the gradient with respect to them, g' = R g; with those it
    reproduces the rigid call's own dL_dpose for a transform that is not the identity."""
    import diff_gaussian_rasterization as D
    import gsr_pose
    P, W, H = 20000, 200, 120
    cam = S.make_camera(W, H)
    sc = S.make_gvol(P, W, H, seed=73, s0=0.02)
    rs = cuda_settings(D, cam, (0.0, 0.0, 0.0), 3, DEV)
    T = gsr_pose.se3_exp(torch.tensor([0.05, -0.03, 0.02, 0.02, -0.04, 0.03], dtype=torch.float64)).float().to(DEV)
    xyz = sc["means3D"].to(DEV).requires_grad_(True)
    xi = torch.zeros(6, device=DEV, requires_grad=True)
    color, radii, depth, alpha = gsr_pose.rasterize_with_pose(
        rs, xi, xyz, T, sc["opacities"].to(DEV), shs=sc["shs"].to(DEV), scales=sc["scales"].to(DEV),
        rotations=sc["rotations"].to(DEV))
    ct = S.cotangents(W, H, seed=74, dense=True)
    ((color * ct["grad_color"].to(DEV)).sum() + (depth * ct["grad_depth"].to(DEV)).sum()).backward()
    R, t = T[:3, :3], T[:3, 3]
    moved = (xyz.detach() @ R.T + t).contiguous()
    g_moved = (xyz.grad @ R.T).contiguous()
    got = _static_pose_gradient(moved, g_moved)
    want = xi.grad
    err = float((got - want).abs().max()) / (float(want.abs().max()) + 1e-30)
    assert err <= 1e-4, (err, got.tolist(), want.tolist())
    # the raw points and gradient give something else
    wrong = _static_pose_gradient(xyz.detach().contiguous(), xyz.grad.contiguous())
    assert float((wrong - want).abs().max()) / (float(want.abs().max()) + 1e-30) > 1e-3


def test_static_scratch_rejects_oversized_and_handles_empty():
    from diff_gaussian_rasterization import _cabi
    out = torch.zeros(6, device=DEV)
    dummy = torch.zeros(3, device=DEV)
    status = _cabi.lib.gsr_pose_gradient_static(MAX_POINTS + 1, dummy.data_ptr(), dummy.data_ptr(), out.data_ptr(), None)
    assert status != 0
    msg = _cabi.lib.gsr_last_error().decode()
    assert "GSR_POSE_STATIC_MAX_POINTS" in msg and str(MAX_POINTS) in msg
    out.fill_(7.0)
    assert _cabi.lib.gsr_pose_gradient_static(0, None, None, out.data_ptr(), torch.cuda.current_stream().cuda_stream) == 0
    torch.cuda.synchronize()
    assert float(out.abs().max()) == 0.0
