"""Where does a CF-3DGS global-step iteration (1.04M Gaussians, 960x540) spend its GPU time? (development aid)"""
import ctypes, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "cf-3dgs_b200")); sys.path.insert(0, ROOT)
import torch, gsr_synth as S, diff_gaussian_rasterization as D, gsr_model, gsr_loss
dev = torch.device("cuda:0"); W, H = 960, 540
cam = S.make_camera(W, H)
sc = S.make_gpix(W, H, seed=3, target_P=130_000)
import numpy as np
xs = torch.cat([sc["means3D"] + 0.01 * i for i in range(8)]).to(dev)
P = xs.shape[0]
shs = torch.cat([sc["shs"]] * 8).to(dev); ls = torch.cat([sc["scales"]] * 8).to(dev).log()
rot = torch.zeros(P, 4, device=dev); rot[:, 0] = 1
opa = torch.full((P, 1), -2.2, device=dev)
m = gsr_model.FusedGaussianModel(xs, shs, opa, ls, rot, dict(xyz=1.6e-4, f_dc=2.5e-3, f_rest=1.25e-4, opacity=0.05, scaling=5e-3, rotation=1e-3))
rs = D.GaussianRasterizationSettings(H, W, cam["tanfovx"], cam["tanfovy"], torch.zeros(3, device=dev), 1.0, cam["viewmatrix"].to(dev),
                                     cam["projmatrix"].to(dev), 3, cam["campos"].to(dev), False, False)
gt = torch.rand(3, H, W, device=dev)
ev = lambda: torch.cuda.Event(enable_timing=True)
acc = {}
def tick(name, e0, e1): acc[name] = acc.get(name, 0.0) + e0.elapsed_time(e1)
lib = D.lib; n_st = lib.gsr_profile_count(); names = [lib.gsr_profile_name(i).decode() for i in range(n_st)]
stage = {n: 0.0 for n in names}
N = 20
for it in range(N + 3):
    e = [ev() for _ in range(6)]
    e[0].record()
    act = m.activated()
    x = m.xyz.detach().requires_grad_(True); sh = m.features.detach().requires_grad_(True)
    o, s_, r = act["opacities"].requires_grad_(True), act["scales"].requires_grad_(True), act["rotations"].requires_grad_(True)
    m2d = torch.zeros_like(x, requires_grad=True)
    e[1].record()
    lib.gsr_profile_enable(1)
    img, radii, _, _ = D.GaussianRasterizer(rs)(x, m2d, o, shs=sh, scales=s_, rotations=r)
    e[2].record()
    loss = gsr_loss.photometric_loss(img, gt, lambda_dssim=0.2, clamp=True)
    e[3].record()
    loss.backward()
    e[4].record()
    buf = (ctypes.c_float * n_st)(); lib.gsr_profile_read(buf, n_st); lib.gsr_profile_enable(0)
    m.step(x.grad, sh.grad, o.grad, s_.grad, r.grad, d_means2D=m2d.grad, radii=radii)
    e[5].record(); torch.cuda.synchronize()
    if it >= 3:
        for k, (a, b) in {"activate+leaves": (0, 1), "raster forward (API)": (1, 2), "loss forward": (2, 3), "backward (loss+raster)": (3, 4), "model step": (4, 5), "total": (0, 5)}.items():
            tick(k, e[a], e[b])
        for n_, v in zip(names, buf[:]): stage[n_] += max(v, 0.0)
print("P", P, {k: round(v / N, 3) for k, v in acc.items()})
print({k: round(v / N, 3) for k, v in stage.items()})
