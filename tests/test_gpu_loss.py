"""GPU: the fused L1 + SSIM loss against the reference's PyTorch formulation (utils/loss_utils.py, trainer/losses.py)."""
import math

import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _window(device, dtype):
    g = torch.tensor([math.exp(-(x - 5) ** 2 / float(2 * 1.5 ** 2)) for x in range(11)])
    g = (g / g.sum()).unsqueeze(1)
    return g.mm(g.t()).float().unsqueeze(0).unsqueeze(0).expand(3, 1, 11, 11).contiguous().to(device=device, dtype=dtype)


def _ref_loss(img, gt, lam, clamp):
    """utils/loss_utils.py:_ssim + l1_loss, combined as trainer/losses.py:113-123."""
    x = img.clamp(0, 1) if clamp else img
    w = _window(x.device, x.dtype)
    x4, y4 = x.unsqueeze(0), gt.unsqueeze(0)
    mu1 = F.conv2d(x4, w, padding=5, groups=3); mu2 = F.conv2d(y4, w, padding=5, groups=3)
    mu1_sq, mu2_sq, mu12 = mu1.pow(2), mu2.pow(2), mu1 * mu2
    s1 = F.conv2d(x4 * x4, w, padding=5, groups=3) - mu1_sq
    s2 = F.conv2d(y4 * y4, w, padding=5, groups=3) - mu2_sq
    s12 = F.conv2d(x4 * y4, w, padding=5, groups=3) - mu12
    C1, C2 = 0.01 ** 2, 0.03 ** 2
    ssim = (((2 * mu12 + C1) * (2 * s12 + C2)) / ((mu1_sq + mu2_sq + C1) * (s1 + s2 + C2))).mean()
    l1 = (x - gt).abs().mean()
    return (1 - lam) * l1 + lam * (1 - ssim), l1, ssim


@pytest.mark.parametrize("shape", [(3, 540, 960), (3, 67, 93), (3, 16, 16)])
@pytest.mark.parametrize("clamp", [False, True])
def test_fused_loss_matches_reference_formulation(shape, clamp):
    import gsr_loss
    g = torch.Generator().manual_seed(5)
    img = (torch.rand(shape, generator=g) * (1.4 if clamp else 1.0) - (0.2 if clamp else 0.0)).to(DEV)
    # smooth-ish target so SSIM is not degenerate
    gt = F.avg_pool2d(torch.rand(1, *shape, generator=g), 5, 1, 2)[0].to(DEV)
    a = img.clone().requires_grad_(True)
    loss, parts = gsr_loss.photometric_loss(a, gt, 0.2, clamp=clamp, return_parts=True)
    (loss * 3.0).backward()
    b = img.double().clone().requires_grad_(True)
    ref, l1, ssim = _ref_loss(b, gt.double(), 0.2, clamp)
    (ref * 3.0).backward()
    assert abs(float(loss) - float(ref)) <= 1e-5 * max(1.0, abs(float(ref)))
    assert abs(float(parts[1]) - float(l1)) <= 1e-5 and abs(float(parts[2]) - float(ssim)) <= 1e-5
    err = float((a.grad.double() - b.grad).abs().max()) / float(b.grad.abs().max())
    assert err <= 1e-4, err


def test_fused_loss_reproduces_the_reference_code_output():
    """Golden vectors made by importing the reference's own trainer/losses.py (tests/golden/make_golden_loss.py)."""
    import glob
    import os
    import numpy as np
    import gsr_loss
    files = sorted(glob.glob(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "loss_*.npz")))
    assert files, "loss fixtures missing"
    for path in files:
        z = np.load(path)
        img = torch.from_numpy(z["render"]).to(DEV).requires_grad_(True)
        gt = torch.from_numpy(z["gt"]).to(DEV)
        loss = gsr_loss.photometric_loss(img, gt, lambda_dssim=float(z["lam"]), clamp=bool(z["clamp"]))
        loss.backward()
        assert abs(float(loss) - float(z["ref_loss"])) <= 1e-5, path
        ref_g = torch.from_numpy(z["ref_grad"]).to(DEV)
        assert float((img.grad.double() - ref_g).abs().max()) <= 1e-4 * float(ref_g.abs().max()), path


@pytest.mark.parametrize("clamp", [False, True])
def test_lambda_zero_is_the_plain_l1_loss(clamp):
    """lambda_dssim = 0 takes the streaming L1 path (no SSIM evaluated): value and gradient against torch."""
    import gsr_loss
    g = torch.Generator().manual_seed(21)
    img = (torch.rand(3, 67, 131, generator=g) * 1.4 - 0.2).to(DEV).requires_grad_(True)
    gt = torch.rand(3, 67, 131, generator=g).to(DEV)
    loss, parts = gsr_loss.photometric_loss(img, gt, 0.0, clamp=clamp, return_parts=True)
    loss.backward()
    ref_in = img.detach().clone().requires_grad_(True)
    x = ref_in.clamp(0, 1) if clamp else ref_in
    ref = (x - gt).abs().mean()
    ref.backward()
    assert abs(float(loss) - float(ref)) <= 1e-6 and float(parts[2]) == 0.0 and abs(float(parts[1]) - float(ref)) <= 1e-6
    assert float((img.grad - ref_in.grad).abs().max()) <= 1e-9
