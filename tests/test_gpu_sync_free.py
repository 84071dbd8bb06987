"""GPU: the host-synchronisation-free forward (gsr_forward_fused) and CUDA-graph capture of a whole fwd+bwd step."""
import warnings

import pytest
import torch

import gsr_synth as S
from helpers import cuda_settings

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


@pytest.fixture(autouse=True)
def _restore_mode():
    import diff_gaussian_rasterization as D
    yield
    D.set_sync_free(False)
    D._SyncFree.capacity.clear(); D._SyncFree.pending.clear(); D._SyncFree.last_status.clear(); D._SyncFree.warned.clear()
    for st in D._SyncFree.device_status.values():
        st.zero_()


def _scene(P=30000, W=320, H=180, seed=81):
    cam = S.make_camera(W, H)
    sc = S.make_gvol(P, W, H, seed=seed)
    return cam, {k: sc[k].to(DEV) for k in ["means3D", "opacities", "shs", "scales", "rotations"]}


def _step(D, rs, t, gc):
    leaves = {k: v.detach().clone().requires_grad_(True) for k, v in t.items()}
    m2d = torch.zeros_like(leaves["means3D"], requires_grad=True)
    color, radii, depth, alpha = D.GaussianRasterizer(rs)(leaves["means3D"], m2d, leaves["opacities"], shs=leaves["shs"],
                                                          scales=leaves["scales"], rotations=leaves["rotations"])
    color.backward(gc)
    return color.detach(), depth.detach(), alpha.detach(), radii, {k: v.grad for k, v in leaves.items()}


def test_sync_free_forward_backward_is_bit_identical_to_the_synchronising_path():
    import diff_gaussian_rasterization as D
    cam, t = _scene()
    rs = cuda_settings(D, cam, (0.1, 0.2, 0.3), 3, DEV)
    gc = S.cotangents(cam["W"], cam["H"], seed=82, dense=False)["grad_color"].to(DEV)
    ref = _step(D, rs, t, gc)
    D.set_sync_free(True, slack=1.3)
    first = _step(D, rs, t, gc)      # learns the capacity through the synchronising path
    assert D._SyncFree.capacity, "capacity was not learnt"
    second = _step(D, rs, t, gc)     # gsr_forward_fused
    R, overflow = D.sync_free_status()
    assert R > 0 and not overflow
    for a, b in zip(ref[:4], second[:4]):
        assert torch.equal(a, b)
    for k in ref[4]:
        # backward accumulates with float atomics: equal up to summation order
        assert float((ref[4][k] - second[4][k]).abs().max()) <= 5e-5 * float(ref[4][k].abs().max())
    assert torch.equal(first[0], ref[0])


def test_sync_free_overflow_is_flagged_and_harmless():
    import diff_gaussian_rasterization as D
    cam, t = _scene(P=20000)
    rs = cuda_settings(D, cam, (0, 0, 0), 3, DEV)
    gc = S.cotangents(cam["W"], cam["H"], seed=83, dense=False)["grad_color"].to(DEV)
    D.set_sync_free(True, capacity=5000)          # far too small on purpose
    out = _step(D, rs, t, gc)
    with warnings.catch_warnings(record=True) as w:
        warnings.simplefilter("always")
        R, overflow = D.sync_free_status()
    assert overflow and R > 5000 and any("exceeded the binning capacity" in str(x.message) for x in w)
    assert bool(torch.isfinite(out[0]).all())
    assert D._SyncFree.capacity[torch.cuda.current_device()] >= R      # raised for the next call
    ok = _step(D, rs, t, gc)
    R2, overflow2 = D.sync_free_status()
    assert not overflow2 and R2 == R
    D.set_sync_free(False)
    ref = _step(D, rs, t, gc)
    assert torch.equal(ok[0], ref[0])


def test_overflow_of_an_earlier_view_or_replay_is_not_lost():
    """One status buffer per device, only ever OR-ed / max-ed into by the kernels: an overflowing render followed by
    renders that fit (another view, later replays of captured graphs) is still reported by the next look."""
    import diff_gaussian_rasterization as D
    cam, big = _scene(P=20000)
    _, small = _scene(P=300, seed=85)
    rs = cuda_settings(D, cam, (0, 0, 0), 3, DEV)
    gc = S.cotangents(cam["W"], cam["H"], seed=83, dense=False)["grad_color"].to(DEV)
    D.set_sync_free(True, capacity=6000)
    D._SyncFree.slack = 0.0                       # keep the capacity where it is: the graphs below are sized for it

    def fwd(t):
        return D.GaussianRasterizer(rs)(t["means3D"], torch.zeros_like(t["means3D"]), t["opacities"], shs=t["shs"],
                                        scales=t["scales"], rotations=t["rotations"])[0]

    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        fwd(small)
    torch.cuda.current_stream().wait_stream(side)
    assert not D.sync_free_status()[1]
    g_big, g_small = torch.cuda.CUDAGraph(), torch.cuda.CUDAGraph()
    with torch.cuda.graph(g_big):
        fwd(big)
    with torch.cuda.graph(g_small):
        fwd(small)
    g_small.replay(); g_small.replay()
    R, overflow = D.sync_free_status()
    assert not overflow and 0 < R <= 6000
    g_big.replay()                                # needs far more than 6000 instances
    g_small.replay(); g_small.replay()            # later renders that fit must not hide it
    with warnings.catch_warnings(record=True):
        warnings.simplefilter("always")
        R, overflow = D.sync_free_status()
    assert overflow and R > 6000
    g_small.replay()
    R, overflow = D.sync_free_status()            # the look cleared the record
    assert not overflow and R <= 6000


def test_whole_step_can_be_captured_in_a_cuda_graph():
    import diff_gaussian_rasterization as D
    cam, t = _scene(P=20000)
    rs = cuda_settings(D, cam, (0, 0, 0), 3, DEV)
    gc = S.cotangents(cam["W"], cam["H"], seed=84, dense=False)["grad_color"].to(DEV)
    ref = _step(D, rs, t, gc)
    D.set_sync_free(True, slack=1.5)
    _step(D, rs, t, gc)                            # learn capacity (synchronising)
    static = {k: v.detach().clone().requires_grad_(True) for k, v in t.items()}
    m2d = torch.zeros_like(static["means3D"], requires_grad=True)
    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):                  # warm-up on a side stream, as torch.cuda.graph asks for
        for _ in range(2):
            c = D.GaussianRasterizer(rs)(static["means3D"], m2d, static["opacities"], shs=static["shs"],
                                         scales=static["scales"], rotations=static["rotations"])[0]
            c.backward(gc)
    torch.cuda.current_stream().wait_stream(side)
    for v in static.values():
        v.grad = None
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        color = D.GaussianRasterizer(rs)(static["means3D"], m2d, static["opacities"], shs=static["shs"],
                                         scales=static["scales"], rotations=static["rotations"])[0]
        color.backward(gc)
    for _ in range(3):
        g.replay()
    torch.cuda.synchronize()
    assert torch.equal(color.detach(), ref[0])
    assert float((static["means3D"].grad - ref[4]["means3D"]).abs().max()) <= 5e-5 * float(ref[4]["means3D"].abs().max())
    # new parameter values flow through the captured graph
    with torch.no_grad():
        static["opacities"].mul_(0.5)
    g.replay()
    torch.cuda.synchronize()
    assert not torch.equal(color.detach(), ref[0])


def test_state_dump_after_sync_free_forward_matches_the_synchronising_path():
    """The binning workspace is sized for the capacity; only the first num_rendered entries are defined."""
    import diff_gaussian_rasterization as D
    from helpers import cuda_forward
    cam = S.second_view(200, 120)
    sc = S.make_gvol(6000, 200, 120, seed=85, s0=0.02)
    ref, _, _ = cuda_forward(sc, cam, (0, 0, 0))
    D.set_sync_free(True, slack=2.0)
    cuda_forward(sc, cam, (0, 0, 0))                     # learns the capacity
    import torch as _t
    leaves = {k: sc[k].to(DEV) for k in ["means3D", "opacities", "shs", "scales", "rotations"]}
    rs = cuda_settings(D, cam, (0, 0, 0), 3, DEV)
    e = _t.Tensor([])
    cap, color, depth, alpha, radii, geom, binning, img = D._forward_impl(
        rs.bg, leaves["means3D"], e, leaves["opacities"], leaves["scales"], leaves["rotations"], 1.0, e, rs.viewmatrix,
        rs.projmatrix, rs.tanfovx, rs.tanfovy, 120, 200, leaves["shs"], 3, rs.campos, False, False)
    R, overflow = D.sync_free_status()
    assert not overflow and R == ref["num_rendered"] and cap > R
    st = D.rasterizer_state(6000, 200, 120, cap, geom, binning, img, actual=R)
    assert _t.equal(st["keys_sorted"], ref["state"]["keys_sorted"])
    assert _t.equal(st["point_list"], ref["state"]["point_list"])
    assert _t.equal(st["ranges"], ref["state"]["ranges"])
    assert _t.equal(color, ref["color"].detach())
