"""GPU: seeded random scenes against the live reference build (oracle/_ref) -- image sizes that are not multiples of the tile,
random cameras, scale ranges from sub-pixel to screen-filling, opacities at the 1/255 and 0.99 thresholds, every SH degree,
points behind and beside the camera.  Same bars as the parity tests: keys / order / ranges / n_contrib bit-exact, images
<= 1e-5 (they come out bit-equal), gradients <= 1e-4 relative."""
import math

import numpy as np
import pytest
import torch

import gsr_synth as S
from helpers import cuda_forward, diag, rel_elementwise, rel_to_max

pytestmark = pytest.mark.gpu
N_CASES = 24


def _case(i):
    rng = np.random.default_rng(1000 + i)
    W = int(rng.integers(17, 520)); H = int(rng.integers(17, 330))
    P = int(rng.integers(1, 30000))
    deg = int(rng.integers(0, 4))
    yaw, pitch = rng.uniform(-25, 25), rng.uniform(-15, 15)
    a, b = math.radians(yaw), math.radians(pitch)
    Ry = np.array([[math.cos(a), 0, math.sin(a)], [0, 1, 0], [-math.sin(a), 0, math.cos(a)]])
    Rx = np.array([[1, 0, 0], [0, math.cos(b), -math.sin(b)], [0, math.sin(b), math.cos(b)]])
    cam = S.make_camera(W, H, fovx_deg=float(rng.uniform(35, 95)), R=Rx @ Ry, t=rng.uniform(-0.4, 0.4, 3))
    # from far below a pixel to splats covering the screen; the strongly anisotropic cases stay below ~100 px of radius:
    # a 1000-pixel needle next to the camera makes the covariance gradients a sum of millions of cancelling float32 terms,
    # and then this build and the reference build both sit ~10 % away from a float64 evaluation, on either side of it
    # (tools/diag_fuzz.py) -- nothing a 1e-4 comparison between the two can say anything about
    aniso = i % 5 == 0
    s0 = float(10 ** rng.uniform(-3.0, -1.9 if aniso else -0.6))
    sc = S.make_gvol(P, W, H, seed=2000 + i, s0=s0, sh_degree=deg)
    g = torch.Generator().manual_seed(3000 + i)
    # opacities that sit on the two thresholds of the blend, exact zeros and ones among them
    pick = torch.rand(P, 1, generator=g)
    special = torch.tensor([0.0, 1.0 / 255.0, 1.0 / 255.0 + 1e-7, 0.99, 0.990001, 1.0])[torch.randint(0, 6, (P, 1), generator=g)]
    sc["opacities"] = torch.where(pick < 0.15, special, sc["opacities"]).contiguous()
    if aniso:        # strongly anisotropic
        sc["scales"] = (sc["scales"] * torch.tensor([1.0, 0.05, 8.0])).contiguous()
    bg = tuple(float(x) for x in rng.uniform(0, 1, 3)) if i % 2 else (0.0, 0.0, 0.0)
    return cam, sc, deg, bg, W, H, P


def _float64_gradients(i, cam, sc, deg, bg, W, H, ct):
    """Gradients of the same loss by float64 autograd through oracle/torch_oracle.py (CPU; a few seconds per case)."""
    from oracle import torch_oracle as TO
    lv = {k: sc[k].double().clone().requires_grad_(True) for k in ["means3D", "opacities", "shs", "scales", "rotations"]}
    o = TO.forward(lv["means3D"], lv["opacities"], shs=lv["shs"], scales=lv["scales"], rotations=lv["rotations"],
                   viewmatrix=cam["viewmatrix"].double(), projmatrix=cam["projmatrix"].double(), campos=cam["campos"].double(),
                   bg=torch.tensor(bg).double(), W=W, H=H, tanfovx=cam["tanfovx"], tanfovy=cam["tanfovy"], sh_degree=deg)
    loss = (o["color"] * ct["grad_color"].double()).sum()
    if i % 3:
        loss = loss + (o["depth"] * ct["grad_depth"].double()).sum() + (o["alpha"] * ct["grad_alpha"].double()).sum()
    loss.backward()
    return {k: v.grad.numpy() for k, v in lv.items()}


@pytest.mark.parametrize("i", range(N_CASES))
def test_random_scene_vs_reference_cuda(i):
    from oracle import ref_cuda as RC
    if not RC.available():
        pytest.skip("oracle/_ref/_dgr_ref.so not present")
    cam, sc, deg, bg, W, H, P = _case(i)
    out, leaves, m2d = cuda_forward(sc, cam, bg, sh_degree=deg)
    dev = out["color"].device
    rst = RC.forward(sc["means3D"].to(dev), sc["opacities"].to(dev), shs=sc["shs"].to(dev), scales=sc["scales"].to(dev),
                     rotations=sc["rotations"].to(dev), viewmatrix=cam["viewmatrix"].to(dev),
                     projmatrix=cam["projmatrix"].to(dev), campos=cam["campos"].to(dev),
                     bg=torch.tensor(bg, dtype=torch.float32, device=dev), W=W, H=H, tanfovx=cam["tanfovx"],
                     tanfovy=cam["tanfovy"], sh_degree=deg)
    ref = RC.parse_state(rst, P, W, H)
    mine = out["state"]
    assert out["num_rendered"] == rst["num_rendered"]
    assert torch.equal(out["radii"], rst["radii"])
    assert torch.equal(mine["keys_sorted"], ref["keys_sorted"])
    assert torch.equal(mine["point_list"], ref["point_list"])
    assert torch.equal(mine["ranges"], ref["ranges"])
    assert torch.equal(mine["n_contrib"], ref["n_contrib"])
    for k in ["color", "depth", "alpha"]:
        assert float((out[k].detach() - rst[k]).abs().max()) <= 1e-5, k
    ct = S.cotangents(W, H, seed=4000 + i, dense=bool(i % 3))
    zero = torch.zeros(1, H, W, device=dev)
    gd = ct["grad_depth"].to(dev) if i % 3 else zero
    ga = ct["grad_alpha"].to(dev) if i % 3 else zero
    loss = (out["color"] * ct["grad_color"].to(dev)).sum()
    if i % 3:
        loss = loss + (out["depth"] * gd).sum() + (out["alpha"] * ga).sum()
    loss.backward()
    gr = RC.backward(rst, ct["grad_color"].to(dev), gd, ga)
    f64 = _float64_gradients(i, cam, sc, deg, bg, W, H, ct) if i % 5 == 0 else None
    for nm, key in [("means3D", "means3D"), ("opacities", "opacities"), ("shs", "sh"), ("scales", "scales"),
                    ("rotations", "rotations")]:
        if float(gr[key].abs().max()) == 0.0:
            assert float(leaves[nm].grad.abs().max()) == 0.0, nm
            continue
        mine, ref = leaves[nm].grad.cpu().numpy(), gr[key].cpu().numpy()
        if f64 is None:
            assert rel_to_max(mine, ref) <= 1e-4, (nm, i)
            # element by element too (entries below a thousandth of the largest are compared on that floor's scale)
            assert rel_elementwise(mine, ref) <= 1e-2, (nm, i)
            continue
        # Needle-shaped Gaussians (160:1) have nearly singular 2-D conics: the covariance chain amplifies float32 rounding
        # about a thousandfold, and this build and the reference build then each sit a few 1e-4 (means) to 1-2 % (scales,
        # rotations) from a float64 evaluation, on either side of it (tools/diag_fuzz.py).  A comparison between the two
        # float32 builds says nothing there; the float64 autograd of the PyTorch restatement is the judge: this build must be
        # in the same place relative to it as the reference build is.  Both distances are float32 rounding noise of
        # different operation orders (measured: 2.7e-3 here against 1.3e-3 there for the rotations of case 0, 8.5e-4 against
        # 9.7e-4 for its scales), so "same place" is a factor of three, plus the 1e-4 of the parity tests.
        truth = f64[nm]
        d_mine, d_ref = rel_to_max(mine, truth), rel_to_max(ref, truth)
        assert d_mine <= 3.0 * d_ref + 1e-4, (nm, i, d_mine, d_ref)
        diag(f"fuzz case {i} {nm}: distance to float64 -- this build {d_mine:.2e}, reference build {d_ref:.2e}")
