"""GPU: the fused pose gradient (row N1) against autograd through a matrix-exponential SE3."""
import numpy as np
import pytest
import torch

import gsr_synth as S
from helpers import cuda_settings

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


@pytest.mark.parametrize("degree,freeze", [(3, False), (0, False), (0, True)],
                         ids=["sh3", "sh0", "sh0_pose_only_geometry_kernel"])
def test_fused_pose_gradient_matches_autograd_through_se3_exp(degree, freeze):
    """freeze: the points need no gradient either, so with SH degree 0 the backward blend runs its geometry-only variant
    (no opacity / colour partials) -- the case of CF-3DGS's pose fit."""
    import diff_gaussian_rasterization as D
    import gsr_pose as GP
    W, H, P = 200, 120, 12000
    cam = S.make_camera(W, H)
    sc = S.make_gvol(P, W, H, seed=61, s0=0.02)
    rs = cuda_settings(D, cam, (0.1, 0.1, 0.1), degree, DEV)
    t = {k: sc[k].to(DEV) for k in ["means3D", "opacities", "shs", "scales", "rotations"]}
    a = np.deg2rad(3.0)
    T = torch.tensor([[np.cos(a), 0, np.sin(a), 0.03], [0, 1, 0, -0.02], [-np.sin(a), 0, np.cos(a), 0.05], [0, 0, 0, 1]],
                     dtype=torch.float32, device=DEV)
    ct = S.cotangents(W, H, seed=62, dense=True)
    gc, gd, ga = ct["grad_color"].to(DEV), ct["grad_depth"].to(DEV), ct["grad_alpha"].to(DEV)

    # fused: dL/dxi straight from the backward kernel
    xi = torch.zeros(6, device=DEV, requires_grad=True)
    xyz = t["means3D"].clone().requires_grad_(not freeze)
    color, radii, depth, alpha = GP.rasterize_with_pose(rs, xi, xyz, T, t["opacities"], shs=t["shs"], scales=t["scales"],
                                                        rotations=t["rotations"])
    ((color * gc).sum() + (depth * gd).sum() + (alpha * ga).sum()).backward()

    # reference: autograd through p' = Exp(xi) T x with the plain operator.  The rotation is applied with the kernels'
    # own operation sequence (three nested fma per coordinate; torch.addcmul is one fma), so that both paths render the
    # same point bits -- the blend backward amplifies a last-ulp change of its inputs far beyond 1e-4, which
    # would otherwise be what this test measures.
    class _FmaTransform(torch.autograd.Function):
        @staticmethod
        def forward(ctx, x, M):
            ctx.save_for_backward(x, M)
            cols = []
            for r in range(3):
                t = torch.addcmul(M[r, 3].expand(x.shape[0]), x[:, 0], M[r, 0])
                t = torch.addcmul(t, x[:, 1], M[r, 1])
                cols.append(torch.addcmul(t, x[:, 2], M[r, 2]))
            return torch.stack(cols, 1).contiguous()

        @staticmethod
        def backward(ctx, g):
            x, M = ctx.saved_tensors
            gM = torch.zeros_like(M)
            gM[:3, :3] = g.t().double().matmul(x.double()).float()
            gM[:3, 3] = g.double().sum(0).float()
            return g @ M[:3, :3], gM

    xi2 = torch.zeros(6, device=DEV, requires_grad=True)
    xyz2 = t["means3D"].clone().requires_grad_(True)
    M = GP.se3_exp(xi2) @ T
    means = _FmaTransform.apply(xyz2, M)
    c2, r2, d2, a2 = D.GaussianRasterizer(rs)(means, torch.zeros_like(means), t["opacities"], shs=t["shs"],
                                              scales=t["scales"], rotations=t["rotations"])
    ((c2 * gc).sum() + (d2 * gd).sum() + (a2 * ga).sum()).backward()

    assert torch.equal(radii, r2) and torch.equal(color, c2), "the fused transform and the fma chain disagree"
    want, got = xi2.grad.double().cpu().numpy(), xi.grad.double().cpu().numpy()
    assert np.abs(want).max() > 0
    assert np.abs(got - want).max() <= 1e-4 * np.abs(want).max(), (got, want)
    # gradient w.r.t. the untransformed points is R^T g
    if freeze:
        assert xyz.grad is None
        return
    assert float((xyz.grad - xyz2.grad).abs().max()) <= 1e-4 * float(xyz2.grad.abs().max())


def test_pose_descent_recovers_a_known_offset():
    """Pose-only fit (CF-3DGS local pose fit, trainer/cf3dgs_trainer.py:262-285) on a G-pix frame pair."""
    import diff_gaussian_rasterization as D
    import gsr_pose as GP
    W, H = 240, 136
    cam = S.make_camera(W, H)
    sc = S.make_gpix(W, H, seed=2, target_P=20000)
    rs = cuda_settings(D, cam, (0, 0, 0), 0, DEV)
    t = {k: sc[k].to(DEV) for k in ["means3D", "opacities", "shs", "scales", "rotations"]}
    t["opacities"] = torch.full_like(t["opacities"], 0.9)
    a = np.deg2rad(0.8)
    T_true = torch.tensor([[np.cos(a), 0, np.sin(a), 0.02], [0, 1, 0, 0.01], [-np.sin(a), 0, np.cos(a), -0.015], [0, 0, 0, 1]],
                          dtype=torch.float32, device=DEV)
    with torch.no_grad():
        tgt = GP.rasterize_with_pose(rs, torch.zeros(6, device=DEV), t["means3D"], T_true, t["opacities"], shs=t["shs"],
                                     scales=t["scales"], rotations=t["rotations"])[0]
    T = torch.eye(4, device=DEV)
    m = torch.zeros(6, device=DEV); v = torch.zeros(6, device=DEV)
    first = None
    for it in range(150):
        xi = torch.zeros(6, device=DEV, requires_grad=True)
        img = GP.rasterize_with_pose(rs, xi, t["means3D"], T, t["opacities"], shs=t["shs"], scales=t["scales"],
                                     rotations=t["rotations"])[0]
        loss = (img - tgt).abs().mean()
        loss.backward()
        first = float(loss) if first is None else first
        g = xi.grad
        m = 0.9 * m + 0.1 * g; v = 0.999 * v + 0.001 * g * g
        step = -2e-3 * (m / (1 - 0.9 ** (it + 1))) / ((v / (1 - 0.999 ** (it + 1))).sqrt() + 1e-8)
        T = (GP.se3_exp(step.double()) @ T.double()).float()
    err_t = float((T[:3, 3] - T_true[:3, 3]).norm())
    err_R = float((T[:3, :3] - T_true[:3, :3]).norm())
    assert float(loss) < 0.25 * first, (first, float(loss))
    assert err_t < 0.01 and err_R < 0.01, (err_t, err_R)


def test_device_pose_step_matches_adam_and_the_se3_exponential():
    """gsr_pose_adam_step against torch.optim.Adam on the twist + the matrix exponential, five steps."""
    import gsr_pose as GP
    from diff_gaussian_rasterization._cabi import check, lib
    g = torch.Generator().manual_seed(11)
    T = torch.eye(4, device=DEV).contiguous()
    state = torch.zeros(13, device=DEV)
    T_ref = torch.eye(4, dtype=torch.float64)
    xi = torch.zeros(6, dtype=torch.float64, requires_grad=True)
    opt = torch.optim.Adam([xi], lr=2e-3, betas=(0.9, 0.999), eps=1e-8)
    for it in range(5):
        grad = torch.randn(6, generator=g) * (10.0 ** float(it - 2))
        check(lib.gsr_pose_adam_step(grad.to(DEV).data_ptr(), state.data_ptr(), T.data_ptr(), 2e-3, 0.9, 0.999, 1e-8,
                                     torch.cuda.current_stream().cuda_stream))
        before = xi.detach().clone()
        xi.grad = grad.double()
        opt.step()
        step = xi.detach() - before                       # what Adam moved the twist by
        T_ref = GP.se3_exp(step) @ T_ref
    torch.cuda.synchronize()
    assert float((T.cpu().double() - T_ref).abs().max()) <= 2e-6
    assert int(state[12]) == 5


def test_pose_optimizer_recovers_a_known_offset_eagerly_and_as_a_graph():
    """The library's pose loop (gsr_pose.PoseOptimizer: no autograd, no PyTorch kernel in the iteration) on the frame pair
    of test_pose_descent_recovers_a_known_offset; the CUDA-graph replay gives the same pose as the eager loop."""
    import diff_gaussian_rasterization as D
    import gsr_pose as GP
    W, H = 240, 136
    cam = S.make_camera(W, H)
    sc = S.make_gpix(W, H, seed=2, target_P=20000)
    rs = cuda_settings(D, cam, (0, 0, 0), 0, DEV)
    t = {k: sc[k].to(DEV) for k in ["means3D", "opacities", "shs", "scales", "rotations"]}
    t["opacities"] = torch.full_like(t["opacities"], 0.9)
    a = np.deg2rad(0.8)
    T_true = torch.tensor([[np.cos(a), 0, np.sin(a), 0.02], [0, 1, 0, 0.01], [-np.sin(a), 0, np.cos(a), -0.015], [0, 0, 0, 1]],
                          dtype=torch.float32, device=DEV)
    with torch.no_grad():
        tgt = GP.rasterize_with_pose(rs, torch.zeros(6, device=DEV), t["means3D"], T_true, t["opacities"], shs=t["shs"],
                                     scales=t["scales"], rotations=t["rotations"])[0].clamp(0, 1).contiguous()
    mk = lambda: GP.PoseOptimizer(rs, t["means3D"], t["opacities"], shs=t["shs"], scales=t["scales"], rotations=t["rotations"],
                                  lr=2e-3)
    eager = mk()
    first = None
    for it in range(150):
        loss = eager.step(tgt)
        first = float(loss) if first is None else first
    assert float(loss) < 0.25 * first, (first, float(loss))
    assert float((eager.T[:3, 3] - T_true[:3, 3]).norm()) < 0.01 and float((eager.T[:3, :3] - T_true[:3, :3]).norm()) < 0.01
    # the same loop as one replayed graph (sync-free forward: no host read of the instance count inside)
    D.set_sync_free(True, slack=2.0)
    try:
        gopt = mk()
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            for _ in range(3):
                gopt.step(tgt)
        torch.cuda.current_stream().wait_stream(side)
        gopt.reset()
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph):
            gopt.step(tgt)
        gopt.reset()
        for it in range(150):
            graph.replay()
        torch.cuda.synchronize()
        R, overflow = D.sync_free_status()
        assert not overflow
        # same arithmetic; the backward's float atomics land in another order, so the poses agree closely, not bit for bit
        assert float((gopt.T - eager.T).abs().max()) < 2e-3
        assert float((gopt.T[:3, 3] - T_true[:3, 3]).norm()) < 0.01
    finally:
        D.set_sync_free(False)
        D._SyncFree.capacity.clear()
