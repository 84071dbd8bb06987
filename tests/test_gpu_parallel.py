"""GPU: the pieces of the batched-global-views loop on one device -- gradients written straight into
slices of a flat buffer, the SH gradient packed to its active bands (dsh_M), the model step applied in parts.  The
collective itself is covered on the CPU (tests/test_dist_gloo.py, world size 2) and on the GPU box by bench.py --gpus N."""
import pytest
import torch

import gsr_synth as S
from helpers import cuda_settings, rel_to_max

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
LRS = dict(xyz=1.6e-4, f_dc=2.5e-3, f_rest=1.25e-4, opacity=0.05, scaling=5e-3, rotation=1e-3)


def _scene(P=20001, W=240, H=136, seed=31):
    cam = S.make_camera(W, H)
    sc = S.make_gvol(P, W, H, seed=seed, s0=0.02)
    return cam, {k: sc[k].to(DEV) for k in ["means3D", "opacities", "shs", "scales", "rotations"]}


@pytest.mark.parametrize("deg", [0, 1, 3])
def test_backward_into_flat_buffer_with_packed_sh_gradient(deg):
    import diff_gaussian_rasterization as D
    import gsr_parallel as GP
    cam, t = _scene()
    P = t["means3D"].shape[0]
    rs = cuda_settings(D, cam, (0.1, 0.2, 0.3), deg, DEV)
    gc = S.cotangents(cam["W"], cam["H"], seed=32, dense=False)["grad_color"].to(DEV)
    e = torch.Tensor([])
    fwd = D._forward_impl(rs.bg, t["means3D"], e, t["opacities"], t["scales"], t["rotations"], 1.0, e, rs.viewmatrix, rs.projmatrix,
                          rs.tanfovx, rs.tanfovy, cam["H"], cam["W"], t["shs"], deg, rs.campos, False, False)
    R, color, depth, alpha, radii, geom, binning, img = fwd
    args = (rs.bg, t["means3D"], radii, e, t["scales"], t["rotations"], 1.0, e, rs.viewmatrix, rs.projmatrix, rs.tanfovx,
            rs.tanfovy, gc, None, None, t["shs"], deg, rs.campos, geom, R, binning, img, alpha, False)
    want = D._backward_impl(*args)                                  # fresh tensors, SH gradient (P, 16, 3)
    M_grad = (deg + 1) ** 2
    lay = GP.flat_layout(P, M_grad)
    flat = torch.full((lay["_total_floats"][0],), float("nan"), device=DEV)
    views = {}
    for name in ("rotations", "means3D", "opacity", "scales", "sh"):
        off, shape = lay[name]
        views[name] = flat[off:off + torch.Size(shape).numel()].view(shape)
    m2d = torch.empty(P, 3, device=DEV)
    got = D._backward_impl(*args, out=dict(means3D=views["means3D"], means2D=m2d, opacity=views["opacity"], sh=views["sh"],
                                           scales=views["scales"], rotations=views["rotations"]), dsh_M=M_grad)
    dmeans2D, _, dopacity, dmeans3D, _, dsh, dscales, drots, _ = got
    assert dmeans3D.data_ptr() == views["means3D"].data_ptr() and dsh.data_ptr() == views["sh"].data_ptr()
    assert bool(torch.isfinite(flat[:11 * P]).all()) and bool(torch.isfinite(views["sh"]).all())   # every slot written
    w2, _, wo, w3, _, wsh, wsc, wr, _ = want
    for a, b in [(dmeans3D, w3), (dopacity, wo), (dscales, wsc), (drots, wr), (dmeans2D, w2), (dsh, wsh[:, :M_grad])]:
        assert rel_to_max(a.cpu().numpy(), b.cpu().numpy()) <= 5e-5      # same kernels, float atomics in another order
    assert float(wsh[:, M_grad:].abs().sum()) == 0.0                       # what packing leaves out is identically zero
    with pytest.raises(D._cabi.GsrError):
        D._backward_impl(*args, dsh_M=M_grad - 1 if M_grad > 1 else 17)    # must hold the active bands, at most M


def test_model_step_in_parts_equals_the_whole_step():
    import gsr_model
    g = torch.Generator().manual_seed(5)
    P, M = 4097, 16
    init = [torch.randn(P, 3, generator=g), torch.randn(P, M, 3, generator=g) * 0.2, torch.randn(P, 1, generator=g),
            torch.randn(P, 3, generator=g) * 0.3 - 3, torch.randn(P, 4, generator=g)]
    whole = gsr_model.FusedGaussianModel(*[t.to(DEV) for t in init], LRS, device_step=True)
    parts = gsr_model.FusedGaussianModel(*[t.to(DEV) for t in init], LRS, device_step=True)
    for it in range(3):
        for M_grad in (16, 4):
            d = dict(means3D=torch.randn(P, 3, generator=g).to(DEV), sh=(torch.randn(P, M_grad, 3, generator=g) * 0.1).to(DEV),
                     opacity=torch.randn(P, 1, generator=g).to(DEV), scales=torch.randn(P, 3, generator=g).to(DEV),
                     rotations=torch.randn(P, 4, generator=g).to(DEV))
            m2d = torch.randn(P, 3, generator=g).to(DEV)
            radii = (torch.rand(P, generator=g) * 10 - 3).clamp(min=0).int().to(DEV)
            full_sh = torch.zeros(P, M, 3, device=DEV)
            full_sh[:, :M_grad] = d["sh"]
            whole.step(d["means3D"], full_sh, d["opacity"], d["scales"], d["rotations"], d_means2D=m2d, radii=radii)
            parts.step_part(1, d, M_grad, tick=True, d_means2D=m2d, radii=radii)
            for row0, rows in ((0, 1500), (1500, 1500), (3000, P - 3000)):
                parts.step_part(2, d, M_grad, row0=row0, rows=rows)
    torch.cuda.synchronize()
    assert int(whole.step_state[0]) == int(parts.step_state[0]) == 6
    for name in ["xyz", "features", "opacity_raw", "log_scales", "rot_raw", "state", "xyz_gradient_accum", "denom", "max_radii2D"]:
        assert torch.equal(getattr(whole, name), getattr(parts, name)), name


def test_batched_views_step_on_one_rank_matches_the_autograd_iteration():
    """World size 1: BatchedViewsStep (direct C-ABI calls, flat buffer, step in parts) against the same iteration written
    with the autograd operator and the whole model step."""
    import diff_gaussian_rasterization as D
    import gsr_loss
    import gsr_model
    import gsr_parallel as GP
    cam, t = _scene(P=15000, seed=33)
    target = torch.rand(3, cam["H"], cam["W"], device=DEV)
    loss_fn = lambda img, gt, window: gsr_loss.photometric_loss(img, gt, 0.2, clamp=True)
    for deg in (0, 3):
        rs = cuda_settings(D, cam, (0, 0, 0), deg, DEV)
        mk = lambda: gsr_model.FusedGaussianModel(t["means3D"], t["shs"], torch.logit(t["opacities"].clamp(1e-3, 1 - 1e-3)),
                                                  t["scales"].log(), t["rotations"], LRS, device_step=True)
        a, b = mk(), mk()
        step = GP.BatchedViewsStep(a)
        for it in range(3):
            la = step(rs, target, loss_fn)
            act = b.activated()
            x = b.xyz.detach().requires_grad_(True); sh = b.features.detach().requires_grad_(True)
            o, sc, r = act["opacities"].requires_grad_(True), act["scales"].requires_grad_(True), act["rotations"].requires_grad_(True)
            m2d = torch.zeros_like(x, requires_grad=True)
            img, radii, _, _ = D.GaussianRasterizer(rs)(x, m2d, o, shs=sh, scales=sc, rotations=r)
            lb = loss_fn(img, target, None)
            lb.backward()
            b.step(x.grad, sh.grad, o.grad, sc.grad, r.grad, d_means2D=m2d.grad, radii=radii)
            assert abs(float(la) - float(lb)) <= 1e-6
        for name in ["xyz", "features", "opacity_raw", "log_scales", "rot_raw"]:
            pa, pb = getattr(a, name), getattr(b, name)
            # Three Adam steps of size ~lr; the two paths differ only in the order of float atomics in the backward.  Adam
            # (eps 1e-15) moves a parameter whose gradient is pure rounding noise by a full +-lr whichever way the noise
            # points, so a few entries may sit up to 3 lr apart; all but a thousandth of them must agree closely.
            d = (pa - pb).abs().reshape(-1)
            lr = max(LRS.values())
            assert float(d.max()) <= 3.01 * 2 * lr, name
            assert float(torch.quantile(d[:1_000_000], 0.999)) <= 2e-2 * lr, name
        assert torch.equal(a.denom, b.denom) and torch.equal(a.max_radii2D, b.max_radii2D)
        assert GP.assert_replicas_identical([a.xyz, a.features])          # trivially true on one rank: exercises the checksum
