"""Golden vectors for the fused photometric loss, produced by IMPORTING the reference's own loss code
(/root/reference/trainer/losses.py: l1_loss, SSIM_V2; pure PyTorch, so it runs in the build container on the CPU):

    python tests/golden/make_golden_loss.py          # writes tests/golden/loss_*.npz

loss = (1 - lambda) * l1_loss(pred, gt) + lambda * (1 - SSIM_V2()(pred, gt)) as in Loss.forward (losses.py:98-136),
pred = clamp(render, 0, 1) as in scene/gaussian_model_cf.py:885, evaluated in float64; stored: render, gt, lambda, the
loss value and d loss / d render."""
import importlib.util
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
REF = os.environ.get("GSR_REFERENCE_ROOT", "/root/reference")


def main():
    spec = importlib.util.spec_from_file_location("ref_losses", os.path.join(REF, "trainer", "losses.py"))
    L = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(L)
    g = torch.Generator().manual_seed(2024)
    cases = {"loss_a": (54, 96, 0.2, True), "loss_b": (37, 51, 0.2, False), "loss_c": (32, 48, 0.5, True)}
    for name, (H, W, lam, clamp) in cases.items():
        yy, xx = torch.meshgrid(torch.linspace(0, 3, H), torch.linspace(0, 4, W), indexing="ij")
        base = torch.stack([torch.sin(xx + c) * torch.cos(yy * (1 + 0.3 * c)) for c in range(3)]) * 0.5 + 0.5
        gt = (base + 0.05 * torch.randn(3, H, W, generator=g)).clamp(0, 1)
        render = base * 1.15 - 0.05 + 0.1 * torch.randn(3, H, W, generator=g)      # leaves [0, 1] in places
        r64 = render.double().requires_grad_(True)
        pred = r64.clamp(0, 1) if clamp else r64
        loss = (1 - lam) * L.l1_loss(pred, gt.double()) + lam * (1 - L.SSIM_V2()(pred, gt.double()))
        loss.backward()
        np.savez_compressed(os.path.join(HERE, name + ".npz"), render=render.numpy(), gt=gt.numpy(),
                            lam=np.float64(lam), clamp=np.int32(clamp), ref_loss=np.float64(loss.item()),
                            ref_grad=r64.grad.numpy())
        print(name, H, W, lam, clamp, loss.item())


if __name__ == "__main__":
    main()
