"""Where does a CF-3DGS pose-fit iteration spend its time? (development aid)"""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "cf-3dgs_b200")); sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "examples"))
import torch, gsr_synth as S, diff_gaussian_rasterization as D, gsr_pose as GP
import cf3dgs_pair as CP
dev = torch.device("cuda:0"); W, H = 960, 540
cam = S.make_camera(W, H); sc = S.make_gpix(W, H, seed=2, target_P=250000)
P = sc["means3D"].shape[0]
rs = D.GaussianRasterizationSettings(H, W, cam["tanfovx"], cam["tanfovy"], torch.zeros(3, device=dev), 1.0, cam["viewmatrix"].to(dev), cam["projmatrix"].to(dev), 0, cam["campos"].to(dev), False, False)
t = {k: sc[k].to(dev) for k in ["means3D", "opacities", "shs", "scales", "rotations"]}
t["opacities"] = torch.full_like(t["opacities"], 0.5)
window = CP._gauss_window(device=dev); gt = torch.rand(3, H, W, device=dev); T = torch.eye(4, device=dev)
def timeit(fn, n=50):
    for _ in range(5): fn()
    torch.cuda.synchronize(); t0 = time.time()
    for _ in range(n): fn()
    torch.cuda.synchronize(); return (time.time() - t0) / n * 1e3
def raster_only():
    xi = torch.zeros(6, device=dev, requires_grad=True)
    img = GP.rasterize_with_pose(rs, xi, t["means3D"], T, t["opacities"], shs=t["shs"], scales=t["scales"], rotations=t["rotations"])[0]
    img.backward(gt)
def full():
    xi = torch.zeros(6, device=dev, requires_grad=True)
    img = GP.rasterize_with_pose(rs, xi, t["means3D"], T, t["opacities"], shs=t["shs"], scales=t["scales"], rotations=t["rotations"])[0].clamp(0, 1)
    CP.photometric_loss(img, gt, window).backward()
img0 = torch.rand(3, H, W, device=dev, requires_grad=True)
def loss_only():
    img0.grad = None
    CP.photometric_loss(img0.clamp(0, 1), gt, window).backward()
def fwd_only():
    with torch.no_grad():
        GP.rasterize_with_pose(rs, torch.zeros(6, device=dev), t["means3D"], T, t["opacities"], shs=t["shs"], scales=t["scales"], rotations=t["rotations"])
print("P", P, "raster fwd+bwd ms %.3f | fwd only %.3f | loss fwd+bwd ms %.3f | full iteration ms %.3f" % (timeit(raster_only), timeit(fwd_only), timeit(loss_only), timeit(full)))
D.lib.gsr_profile_enable(1)
import ctypes
raster_only(); n = D.lib.gsr_profile_count(); buf = (ctypes.c_float * n)(); D.lib.gsr_profile_read(buf, n)
print({D.lib.gsr_profile_name(i).decode(): round(buf[i], 3) for i in range(n)})
