"""Short fwd+bwd loop of the bench workload for ncu (kept tiny: ncu replays every kernel ~40x)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "cf-3dgs_b200")); sys.path.insert(0, ROOT)
import torch
import gsr_synth as S
import diff_gaussian_rasterization as D

P = int(os.environ.get("GSR_P", 1_000_000)); W = int(os.environ.get("GSR_W", 1920)); H = int(os.environ.get("GSR_H", 1080))
steps = int(sys.argv[1]) if len(sys.argv) > 1 else 3
dev = torch.device("cuda:0")
cam = S.make_camera(W, H)
sc = S.make_gvol(P, W, H, seed=5)
params = {k: sc[k].to(dev) for k in ["means3D", "opacities", "shs", "scales", "rotations"]}
gc = S.cotangents(W, H, seed=1, dense=False)["grad_color"].to(dev)
rs = D.GaussianRasterizationSettings(image_height=H, image_width=W, tanfovx=cam["tanfovx"], tanfovy=cam["tanfovy"],
                                     bg=torch.zeros(3, device=dev), scale_modifier=1.0, viewmatrix=cam["viewmatrix"].to(dev),
                                     projmatrix=cam["projmatrix"].to(dev), sh_degree=3, campos=cam["campos"].to(dev),
                                     prefiltered=False, debug=False)
for _ in range(steps):
    leaves = {k: v.detach().requires_grad_(True) for k, v in params.items()}
    m2d = torch.zeros_like(leaves["means3D"], requires_grad=True)
    color = D.GaussianRasterizer(rs)(leaves["means3D"], m2d, leaves["opacities"], shs=leaves["shs"], scales=leaves["scales"],
                                     rotations=leaves["rotations"])[0]
    color.backward(gc)
torch.cuda.synchronize()
print("ok")
