"""GPU: the one timing comparison of the suite (fused loss against its PyTorch formulation), kept apart from the parity
files and ordered after them (tests/conftest.py) so that a noisy box cannot stop a `-x` run before parity has run."""
import pytest
import torch

from test_gpu_loss import _ref_loss

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def test_fused_loss_is_faster_than_the_torch_formulation():
    import gsr_loss
    img = torch.rand(3, 540, 960, device=DEV)
    gt = torch.rand(3, 540, 960, device=DEV)

    def run(fn, n=20):
        for _ in range(3):
            fn()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(n):
            fn()
        e1.record(); torch.cuda.synchronize()
        return e0.elapsed_time(e1) / n

    def fused():
        a = img.clone().requires_grad_(True)
        gsr_loss.photometric_loss(a, gt, 0.2, clamp=True).backward()

    def plain():
        a = img.clone().requires_grad_(True)
        _ref_loss(a, gt, 0.2, True)[0].backward()

    tf, tp = run(fused), run(plain)
    print(f"fused {tf:.3f} ms  torch {tp:.3f} ms")
    assert tf < tp
