"""GPU: fused activation / Adam / densification-statistics kernels against torch.optim.Adam + autograd through the
reference's activations (scene/gaussian_model_cf.py:41-56,242-268,716-719)."""
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
LRS = dict(xyz=1.6e-4, f_dc=2.5e-3, f_rest=1.25e-4, opacity=0.05, scaling=5e-3, rotation=1e-3)


@pytest.mark.parametrize("P", [5000, 4999, 1])     # odd counts: every sub-array of the moment buffer then starts off 16 bytes
def test_fused_model_step_matches_torch_adam_through_the_activations(P):
    import gsr_model
    g = torch.Generator().manual_seed(9)
    M = 16
    xyz = torch.randn(P, 3, generator=g).to(DEV)
    feat = (0.3 * torch.randn(P, M, 3, generator=g)).to(DEV)
    opa = torch.randn(P, 1, generator=g).to(DEV)
    ls = (torch.randn(P, 3, generator=g) * 0.3 - 3).to(DEV)
    rot = torch.randn(P, 4, generator=g).to(DEV)
    model = gsr_model.FusedGaussianModel(xyz, feat, opa, ls, rot, LRS)
    # torch reference: six parameter groups, Adam eps 1e-15, autograd through sigmoid / exp / normalize / cat
    t = dict(xyz=xyz.clone().requires_grad_(True), f_dc=feat[:, :1].clone().requires_grad_(True),
             f_rest=feat[:, 1:].clone().requires_grad_(True), opa=opa.clone().requires_grad_(True),
             ls=ls.clone().requires_grad_(True), rot=rot.clone().requires_grad_(True))
    opt = torch.optim.Adam([{"params": [t["xyz"]], "lr": LRS["xyz"]}, {"params": [t["f_dc"]], "lr": LRS["f_dc"]},
                            {"params": [t["f_rest"]], "lr": LRS["f_rest"]}, {"params": [t["opa"]], "lr": LRS["opacity"]},
                            {"params": [t["ls"]], "lr": LRS["scaling"]}, {"params": [t["rot"]], "lr": LRS["rotation"]}], eps=1e-15)
    accum = torch.zeros(P, 1, device=DEV); denom = torch.zeros(P, 1, device=DEV); maxr = torch.zeros(P, device=DEV)
    for it in range(5):
        gg = torch.Generator().manual_seed(100 + it)
        # "rasterizer" gradients w.r.t. the activated tensors
        d_xyz = torch.randn(P, 3, generator=gg).to(DEV); d_f = torch.randn(P, M, 3, generator=gg).to(DEV)
        d_o = torch.randn(P, 1, generator=gg).to(DEV); d_s = torch.randn(P, 3, generator=gg).to(DEV)
        d_r = torch.randn(P, 4, generator=gg).to(DEV)
        d_m2 = torch.randn(P, 3, generator=gg).to(DEV)
        radii = (torch.rand(P, generator=gg) * 20 - 5).clamp(min=0).int().to(DEV)
        act = model.activated()
        o_ref, s_ref, r_ref = torch.sigmoid(t["opa"]), torch.exp(t["ls"]), F.normalize(t["rot"])
        for a, b in [(act["opacities"], o_ref), (act["scales"], s_ref), (act["rotations"], r_ref)]:
            assert float((a - b).abs().max()) <= 2e-6 * max(1.0, float(b.abs().max()))
        feats = torch.cat([t["f_dc"], t["f_rest"]], 1)
        loss = (t["xyz"] * d_xyz).sum() + (feats * d_f).sum() + (o_ref * d_o).sum() + (s_ref * d_s).sum() + (r_ref * d_r).sum()
        loss.backward()
        opt.step(); opt.zero_grad(set_to_none=True)
        vis = radii > 0
        accum[vis] += torch.norm(d_m2[vis, :2], dim=-1, keepdim=True); denom[vis] += 1
        maxr[vis] = torch.max(maxr[vis], radii[vis].float())
        model.step(d_xyz, d_f, d_o, d_s, d_r, d_means2D=d_m2, radii=radii)
    pairs = [(model.xyz, t["xyz"]), (model.features, torch.cat([t["f_dc"], t["f_rest"]], 1)), (model.opacity_raw, t["opa"]),
             (model.log_scales, t["ls"]), (model.rot_raw, t["rot"])]
    for a, b in pairs:
        # five Adam steps of size ~lr each: agreement far below one step
        assert float((a - b.detach()).abs().max()) <= 2e-5 * max(1e-2, float((b.detach() - 0).abs().max())) + 1e-6
    assert float((model.xyz_gradient_accum - accum[:, 0]).abs().max()) <= 1e-4
    assert torch.equal(model.denom, denom[:, 0]) and torch.equal(model.max_radii2D, maxr)


def test_device_side_step_counter_matches_the_host_one_and_replays_in_a_graph():
    """gsr_gaussian_adam_step_dev keeps the step number on the device: same updates as the host-counted call, also
    when the step is captured once in a CUDA graph and replayed."""
    import gsr_model
    P, M = 3000, 16
    g = torch.Generator().manual_seed(4)
    init = [torch.randn(P, 3, generator=g), torch.randn(P, M, 3, generator=g) * 0.2, torch.randn(P, 1, generator=g),
            torch.randn(P, 3, generator=g) * 0.3 - 3, torch.randn(P, 4, generator=g)]
    lrs = dict(xyz=1e-3, f_dc=2e-3, f_rest=1e-4, opacity=0.05, scaling=5e-3, rotation=1e-3)
    host = gsr_model.FusedGaussianModel(*[t.to(DEV) for t in init], lrs)
    dev_eager = gsr_model.FusedGaussianModel(*[t.to(DEV) for t in init], lrs, device_step=True)
    dev_graph = gsr_model.FusedGaussianModel(*[t.to(DEV) for t in init], lrs, device_step=True)
    grads = [torch.randn(P, 3, generator=g).to(DEV), (torch.randn(P, M, 3, generator=g) * 0.1).to(DEV),
             torch.randn(P, 1, generator=g).to(DEV), torch.randn(P, 3, generator=g).to(DEV), torch.randn(P, 4, generator=g).to(DEV)]
    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        dev_graph.step(*grads)                      # step 1 eagerly (warm-up outside the capture)
    torch.cuda.current_stream().wait_stream(side)
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph):
        dev_graph.step(*grads)                      # step 2 is captured and executed by the first replay
    for _ in range(6):
        host.step(*grads); dev_eager.step(*grads)
    for _ in range(5):
        graph.replay()                               # steps 2..6
    torch.cuda.synchronize()
    assert int(dev_eager.step_state[0]) == 6 and int(dev_graph.step_state[0]) == 6
    for name in ["xyz", "features", "opacity_raw", "log_scales", "rot_raw", "state"]:
        a, b, c = getattr(host, name), getattr(dev_eager, name), getattr(dev_graph, name)
        assert torch.equal(a, b), name
        assert torch.equal(a, c), name


def test_skipping_never_active_sh_bands_changes_nothing():
    """step(..., sh_degree=d) leaves the coefficients of bands that never had a gradient alone; the result is bit-identical
    to the full step on the same (zero above the active bands) gradient, also after the degree has grown."""
    import gsr_model
    P, M = 3001, 16
    g = torch.Generator().manual_seed(6)
    init = [torch.randn(P, 3, generator=g), torch.randn(P, M, 3, generator=g) * 0.2, torch.randn(P, 1, generator=g),
            torch.randn(P, 3, generator=g) * 0.3 - 3, torch.randn(P, 4, generator=g)]
    full = gsr_model.FusedGaussianModel(*[t.to(DEV) for t in init], LRS, device_step=True)
    lean = gsr_model.FusedGaussianModel(*[t.to(DEV) for t in init], LRS, device_step=True)
    for it, deg in enumerate([0, 0, 1, 1, 0, 3]):
        n = (deg + 1) ** 2
        d_f = torch.zeros(P, M, 3, device=DEV)
        d_f[:, :n] = (torch.randn(P, n, 3, generator=g) * 0.1).to(DEV)
        rest = [torch.randn(P, 3, generator=g).to(DEV), torch.randn(P, 1, generator=g).to(DEV),
                torch.randn(P, 3, generator=g).to(DEV), torch.randn(P, 4, generator=g).to(DEV)]
        full.step(rest[0], d_f, rest[1], rest[2], rest[3])
        lean.step(rest[0], d_f, rest[1], rest[2], rest[3], sh_degree=deg)
        assert lean.sh_ever == max((d + 1) ** 2 for d in [0, 0, 1, 1, 0, 3][:it + 1])
    torch.cuda.synchronize()
    for name in ["xyz", "features", "opacity_raw", "log_scales", "rot_raw", "state"]:
        assert torch.equal(getattr(full, name), getattr(lean, name)), name


@pytest.mark.parametrize("M,M_grad,M_upd", [(16, 16, 16), (16, 1, 1), (16, 4, 9), (4, 4, 4), (16, 9, 16), (9, 4, 9), (1, 1, 1)])
def test_shared_step_on_plain_addresses_equals_the_step_in_parts(M, M_grad, M_upd):
    """gsr_gaussian_adam_step_shared with multicast = 0 (plain loads and stores) over two row ranges of a padded model
    gives what gsr_gaussian_adam_step_part gives on the same arrays: the multi-GPU form differs from it only in how the
    summed gradient is read and where the new parameters are written.  Float4 and scalar SH packings."""
    import ctypes
    from diff_gaussian_rasterization._cabi import check, lib
    Pp = 4 * 613
    g = torch.Generator().manual_seed(3)
    mk = lambda *s, sc=1.0: (sc * torch.randn(*s, generator=g)).to(DEV)
    par = dict(xyz=mk(Pp, 3), feat=mk(Pp, M, 3, sc=0.3), opa=mk(Pp, 1), ls=mk(Pp, 3, sc=0.3) - 3, rot=mk(Pp, 4))
    par["rot"][7] = 0                                       # a degenerate quaternion takes the eps branch
    lr = (ctypes.c_float * 6)(1.6e-4, 2.5e-3, 1.25e-4, 0.05, 5e-3, 1e-3)
    n_state = lib.gsr_gaussian_adam_state_floats(Pp, M)
    a = {k: v.clone() for k, v in par.items()}
    b = {k: v.clone() for k, v in par.items()}
    st_a = torch.zeros(n_state, device=DEV); st_b = torch.zeros(n_state, device=DEV)
    ss_a = torch.zeros(2, dtype=torch.int64, device=DEV); ss_b = torch.zeros(2, dtype=torch.int64, device=DEV)
    s = torch.cuda.current_stream().cuda_stream
    for it in range(3):
        gr = dict(xyz=mk(Pp, 3), feat=mk(Pp, M_grad, 3), opa=mk(Pp, 1), ls=mk(Pp, 3), rot=mk(Pp, 4))
        check(lib.gsr_gaussian_adam_step_part(
            Pp, M, M_grad, M_upd, 3, 0, Pp, 1, ss_a.data_ptr(), 0.9, 0.999, 1e-15, lr, a["xyz"].data_ptr(), a["feat"].data_ptr(),
            a["opa"].data_ptr(), a["ls"].data_ptr(), a["rot"].data_ptr(), gr["xyz"].data_ptr(), gr["feat"].data_ptr(),
            gr["opa"].data_ptr(), gr["ls"].data_ptr(), gr["rot"].data_ptr(), st_a.data_ptr(), None, None, None, None, None, s))
        split = 4 * 200
        for i, (row0, rows) in enumerate([(0, split), (split, Pp - split)]):
            ptrs = [b[k].data_ptr() for k in ("xyz", "feat", "opa", "ls", "rot")]
            check(lib.gsr_gaussian_adam_step_shared(
                Pp, M, M_grad, M_upd, row0, rows, 0, int(i == 0), ss_b.data_ptr(), 0.9, 0.999, 1e-15, lr, *ptrs, *ptrs,
                *[gr[k].data_ptr() for k in ("xyz", "feat", "opa", "ls", "rot")], st_b.data_ptr(), s))
    torch.cuda.synchronize()
    assert torch.equal(ss_a, ss_b)
    for k in par:
        assert torch.allclose(a[k], b[k], rtol=1e-6, atol=1e-7), (k, float((a[k] - b[k]).abs().max()))
    # (moments: b1 m + (1 - b1) g cancels for some elements, so the bar is absolute; the two kernels may contract differently)
    assert torch.allclose(st_a, st_b, rtol=1e-6, atol=1e-6), float((st_a - st_b).abs().max())
    # arguments that cannot be moved as float4 are refused
    ptrs = [b[k].data_ptr() for k in ("xyz", "feat", "opa", "ls", "rot")]
    rc = lib.gsr_gaussian_adam_step_shared(Pp, M, M_grad, M_upd, 2, 8, 0, 0, ss_b.data_ptr(), 0.9, 0.999, 1e-15, lr, *ptrs, *ptrs,
                                           *ptrs, st_b.data_ptr(), s)
    assert rc != 0


def test_densify_stats_alone_match_the_reference_update():
    from diff_gaussian_rasterization._cabi import check, lib
    P = 4321
    g = torch.Generator().manual_seed(5)
    m2 = torch.randn(P, 3, generator=g).to(DEV)
    radii = (torch.rand(P, generator=g) * 20 - 5).clamp(min=0).int().to(DEV)
    accum = torch.rand(P, generator=g).to(DEV); denom = torch.ones(P, device=DEV); maxr = torch.full((P,), 3.0, device=DEV)
    want_a, want_d, want_m = accum.clone(), denom.clone(), maxr.clone()
    vis = radii > 0
    want_a[vis] += torch.norm(m2[vis, :2], dim=-1); want_d[vis] += 1; want_m[vis] = torch.max(want_m[vis], radii[vis].float())
    check(lib.gsr_densify_stats(P, m2.data_ptr(), radii.data_ptr(), accum.data_ptr(), denom.data_ptr(), maxr.data_ptr(),
                                torch.cuda.current_stream().cuda_stream))
    assert float((accum - want_a).abs().max()) <= 1e-6 and torch.equal(denom, want_d) and torch.equal(maxr, want_m)


def test_dense_coefficient_0_plane_at_sh_degree_0_equals_stepping_the_full_tensor():
    """shs(0) hands out a dense (P,1,3) plane and step() updates that; a model stepped with the full (P,16,3) gradient
    (zeros above coefficient 0) must end up bit-identical -- at degree 0, after the switch to degree 1, and through the
    `features` property at any time."""
    import gsr_model
    P, M = 4999, 16
    g = torch.Generator().manual_seed(11)
    mk = lambda *s, sc=1.0: (sc * torch.randn(*s, generator=g)).to(DEV)
    init = (mk(P, 3), mk(P, M, 3, sc=0.3), mk(P, 1), mk(P, 3, sc=0.3) - 3, mk(P, 4))
    A = gsr_model.FusedGaussianModel(*init, LRS, device_step=True)
    B = gsr_model.FusedGaussianModel(*init, LRS, device_step=True)
    for it in range(4):
        d = dict(x=mk(P, 3), f0=mk(P, 1, 3), o=mk(P, 1), s=mk(P, 3), r=mk(P, 4))
        sh = A.shs(0)
        assert sh.shape == (P, 1, 3) and sh.is_contiguous() and A.dc_mode
        A.step(d["x"], d["f0"], d["o"], d["s"], d["r"], sh_degree=0)
        full = torch.zeros(P, M, 3, device=DEV)
        full[:, :1] = d["f0"]
        B.step(d["x"], full, d["o"], d["s"], d["r"], sh_degree=0)
        assert torch.equal(A.features, B.features)            # the property brings coefficient 0 in
    for it in range(2):
        d = dict(x=mk(P, 3), f=mk(P, M, 3), o=mk(P, 1), s=mk(P, 3), r=mk(P, 4))
        d["f"][:, 4:] = 0
        assert A.shs(1).shape == (P, M, 3) and not A.dc_mode
        for mdl in (A, B):
            mdl.step(d["x"], d["f"], d["o"], d["s"], d["r"], sh_degree=1)
    for a, b in zip((A.xyz, A.features, A.opacity_raw, A.log_scales, A.rot_raw),
                    (B.xyz, B.features, B.opacity_raw, B.log_scales, B.rot_raw)):
        assert torch.equal(a, b)
    assert torch.equal(A.state, B.state)
    # a (P,1,3) gradient for a model that is not in dc_mode is a caller's mistake
    with pytest.raises(ValueError):
        B.step(d["x"], d["f"][:, :1].contiguous(), d["o"], d["s"], d["r"], sh_degree=0)


def test_global_loop_with_a_rising_sh_degree_matches_the_loop_that_never_uses_the_dc_plane():
    """examples/cf3dgs_global.py with --sh-up-every: degree 0 (dense coefficient-0 plane), then 1 and 2 (full tensor).  The
    same loop rendering from the full tensor throughout (--dc-plane off) must give the same losses: the plane changes where
    coefficient 0 is kept, not what is computed."""
    import os, sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    sys.path.insert(0, os.path.join(root, "examples"))
    import cf3dgs_global as CG
    res = {}
    for mode in ("on", "off"):
        args = CG.default_config(frames=4, points_per_frame=20_000, iters=45, width=320, height=180, sh_degree=2,
                                 sh_up_every=15, warmup=3, dc_plane=mode)
        res[mode] = CG.run_global(args, torch.device(DEV))
    assert res["on"]["last_loss"] < 0.6 * res["on"]["first_loss"]
    assert abs(res["on"]["last_loss"] - res["off"]["last_loss"]) <= 1e-3 * abs(res["off"]["last_loss"])
