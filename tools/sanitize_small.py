"""Small fwd+bwd (all input variants) + loss + pose-static for compute-sanitizer runs."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "cf-3dgs_b200")); sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch, gsr_synth as S
from helpers import add_alt_inputs, cuda_forward
import diff_gaussian_rasterization as D, gsr_loss
for (P, W, H, view, mode, cov, deg) in [(3000, 100, 60, 0, "sh", "scale_rot", 3), (3000, 99, 61, 1, "colors", "cov3D", 0), (1, 33, 17, 0, "sh", "scale_rot", 1)]:
    cam = S.make_camera(W, H) if view == 0 else S.second_view(W, H)
    sc = add_alt_inputs(S.make_gvol(P, W, H, seed=3, s0=0.05, sh_degree=deg), cam)
    out, leaves, m2d = cuda_forward(sc, cam, (0.1, 0.2, 0.3), sh_degree=deg, mode=mode, cov_mode=cov)
    ct = S.cotangents(W, H, seed=4, dense=True)
    loss = gsr_loss.photometric_loss(out["color"], ct["grad_color"].to("cuda").abs().clamp(0, 1), 0.2, clamp=True)
    (loss + (out["depth"] * ct["grad_depth"].to("cuda")).sum() + (out["alpha"] * ct["grad_alpha"].to("cuda")).sum()).backward()
D.set_sync_free(True, 1.5)
for _ in range(2):
    out, leaves, m2d = cuda_forward(sc, cam, (0, 0, 0), sh_degree=1)
    out["color"].sum().backward()
torch.cuda.synchronize(); print("sanitize workload ok")
