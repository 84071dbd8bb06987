"""Shared test helpers: scene construction and calls into the three implementations."""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (os.path.join(ROOT, "cf-3dgs_b200"), ROOT):
    if p not in sys.path:
        sys.path.insert(0, p)

import gsr_synth as S  # noqa: E402

PARAMS = ["means3D", "opacities", "shs", "scales", "rotations"]


def np_cam(cam):
    return dict(viewmatrix=cam["viewmatrix"].numpy(), projmatrix=cam["projmatrix"].numpy(),
                campos=cam["campos"].numpy(), W=cam["W"], H=cam["H"], tanfovx=cam["tanfovx"], tanfovy=cam["tanfovy"])


def oracle_forward(sc, cam, bg, sh_degree=3, mode="sh", cov_mode="scale_rot", scale_modifier=1.0):
    from oracle import oracle as O
    kw = {}
    if mode == "sh":
        kw["shs"] = sc["shs"].numpy()
    else:
        kw["colors_precomp"] = sc["colors"].numpy()
    if cov_mode == "scale_rot":
        kw["scales"] = sc["scales"].numpy(); kw["rotations"] = sc["rotations"].numpy()
    else:
        kw["cov3D_precomp"] = sc["cov3D"].numpy()
    return O.forward(sc["means3D"].numpy(), sc["opacities"].numpy(), bg=np.asarray(bg, np.float32), sh_degree=sh_degree,
                     scale_modifier=scale_modifier, **kw, **np_cam(cam))


def cuda_settings(D, cam, bg, sh_degree, dev, debug=False, scale_modifier=1.0, prefiltered=False):
    return D.GaussianRasterizationSettings(
        image_height=cam["H"], image_width=cam["W"], tanfovx=cam["tanfovx"], tanfovy=cam["tanfovy"],
        bg=torch.as_tensor(bg, dtype=torch.float32).to(dev), scale_modifier=scale_modifier,
        viewmatrix=cam["viewmatrix"].to(dev), projmatrix=cam["projmatrix"].to(dev), sh_degree=sh_degree,
        campos=cam["campos"].to(dev), prefiltered=prefiltered, debug=debug)


def cuda_forward(sc, cam, bg, sh_degree=3, mode="sh", cov_mode="scale_rot", dev="cuda:0", debug=False,
                 scale_modifier=1.0):
    """Runs the product path through the drop-in API.  Returns (outputs dict, leaves dict, means2D holder)."""
    import diff_gaussian_rasterization as D
    dev = torch.device(dev)
    leaves = {}
    kw = {}
    leaves["means3D"] = sc["means3D"].to(dev).clone().requires_grad_(True)
    leaves["opacities"] = sc["opacities"].to(dev).clone().requires_grad_(True)
    if mode == "sh":
        leaves["shs"] = sc["shs"].to(dev).clone().requires_grad_(True); kw["shs"] = leaves["shs"]
    else:
        leaves["colors"] = sc["colors"].to(dev).clone().requires_grad_(True); kw["colors_precomp"] = leaves["colors"]
    if cov_mode == "scale_rot":
        leaves["scales"] = sc["scales"].to(dev).clone().requires_grad_(True)
        leaves["rotations"] = sc["rotations"].to(dev).clone().requires_grad_(True)
        kw["scales"] = leaves["scales"]; kw["rotations"] = leaves["rotations"]
    else:
        leaves["cov3D"] = sc["cov3D"].to(dev).clone().requires_grad_(True); kw["cov3D_precomp"] = leaves["cov3D"]
    m2d = torch.zeros_like(leaves["means3D"], requires_grad=True)
    rs = cuda_settings(D, cam, bg, sh_degree, dev, debug=debug, scale_modifier=scale_modifier)
    color, radii, depth, alpha = D.GaussianRasterizer(rs)(leaves["means3D"], m2d, leaves["opacities"], **kw)
    fn = color.grad_fn
    P = sc["means3D"].shape[0]
    state = D.rasterizer_state(P, cam["W"], cam["H"], fn.num_rendered, fn.saved_tensors[7], fn.saved_tensors[8],
                               fn.saved_tensors[9])
    out = dict(color=color, radii=radii, depth=depth, alpha=alpha, num_rendered=fn.num_rendered, state=state)
    return out, leaves, m2d


def add_alt_inputs(sc, cam):
    """colors_precomp and cov3D_precomp variants of a G-vol scene (CPU tensors)."""
    from oracle import oracle as O
    st = O.forward(sc["means3D"].numpy(), sc["opacities"].numpy(), shs=sc["shs"].numpy(), scales=sc["scales"].numpy(),
                   rotations=sc["rotations"].numpy(), bg=np.zeros(3, np.float32), sh_degree=sc["sh_degree"], **np_cam(cam))
    g = torch.Generator().manual_seed(99)
    sc = dict(sc)
    sc["colors"] = torch.rand(sc["means3D"].shape[0], 3, generator=g)
    cov = torch.from_numpy(st["cov3D"]).clone()
    # Gaussians culled before cov3D was computed have zeros there: give them a valid covariance
    bad = cov.abs().sum(1) == 0
    cov[bad] = torch.tensor([1e-4, 0, 0, 1e-4, 0, 1e-4])
    sc["cov3D"] = cov
    return sc


def rel_to_max(a, b):
    a = np.asarray(a, np.float64).reshape(-1); b = np.asarray(b, np.float64).reshape(-1)
    return float(np.abs(a - b).max() / (np.abs(b).max() + 1e-30))


def rel_elementwise(a, b, floor_frac=1e-3):
    """Largest element-wise relative difference |a - b| / max(|b|, floor), floor = floor_frac x the tensor's largest
    entry: entries below a thousandth of the largest one are compared on the floor's scale (they are sums of float32
    atomics whose own run-to-run spread is of that size), everything above it gets a true relative comparison."""
    a = np.asarray(a, np.float64).reshape(-1); b = np.asarray(b, np.float64).reshape(-1)
    floor = floor_frac * (np.abs(b).max() + 1e-30)
    return float((np.abs(a - b) / np.maximum(np.abs(b), floor)).max())


def diag(msg: str) -> None:
    """Diagnostics that are reported, not asserted (e.g. how many image values are bit-equal to the reference build):
    printed (visible with -s / on failure) and appended to gpurun_out/diagnostics.log, which travels back from the box."""
    print("[diag] " + msg)
    try:
        d = os.path.join(ROOT, "gpurun_out")
        os.makedirs(d, exist_ok=True)
        with open(os.path.join(d, "diagnostics.log"), "a") as f:
            f.write(msg + "\n")
    except OSError:
        pass


def bit_report(name: str, mine, ref) -> None:
    """How far two float32 images are from bit-equality (diagnostic)."""
    a = mine.detach().contiguous().view(torch.int32).long().reshape(-1)
    b = ref.detach().contiguous().view(torch.int32).long().reshape(-1)
    ne = a != b
    n = int(ne.sum())
    worst = int((a - b).abs()[ne].max()) if n else 0
    diag(f"{name}: {n} of {a.numel()} values differ in bits (largest distance {worst} ulp, "
         f"max abs {float((mine.detach() - ref.detach()).abs().max()):.3e})")
