import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "cf-3dgs_b200")); sys.path.insert(0, ROOT)
import torch, gsr_synth as S, diff_gaussian_rasterization as D
dev = torch.device("cuda:0"); P, W, H = 1000000, 1920, 1080
cam = S.make_camera(W, H); sc = S.make_gvol(P, W, H, seed=5)
params = {k: sc[k].to(dev) for k in ["means3D", "opacities", "shs", "scales", "rotations"]}
rs = D.GaussianRasterizationSettings(H, W, cam["tanfovx"], cam["tanfovy"], torch.zeros(3, device=dev), 1.0, cam["viewmatrix"].to(dev), cam["projmatrix"].to(dev), 3, cam["campos"].to(dev), False, False)
tgt = torch.rand(3, H, W, device=dev)
leaves = {k: v.detach().clone().requires_grad_(True) for k, v in params.items()}
out = {}
def compute():
    m2d = torch.zeros_like(leaves["means3D"], requires_grad=True)
    color = D.GaussianRasterizer(rs)(leaves["means3D"], m2d, leaves["opacities"], shs=leaves["shs"], scales=leaves["scales"], rotations=leaves["rotations"])[0]
    loss = (color - tgt).abs().mean(); loss.backward(); out["color"] = color.detach(); out["loss"] = loss.detach()
def t(fn, n=10):
    fn(); torch.cuda.synchronize(); e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): fn()
    e1.record(); torch.cuda.synchronize(); return e0.elapsed_time(e1) / n
print("eager sync ms", t(compute)); ref = out["color"].clone(); gref = leaves["means3D"].grad.clone() / 11
D.set_sync_free(True, 1.5); compute()
print("eager sync-free ms", t(compute), D.sync_free_status())
side = torch.cuda.Stream(); side.wait_stream(torch.cuda.current_stream())
with torch.cuda.stream(side): compute(); compute()
torch.cuda.current_stream().wait_stream(side)
for v in leaves.values(): v.grad = None
g = torch.cuda.CUDAGraph()
with torch.cuda.graph(g): compute()
print("graph ms", t(g.replay), "color equal", torch.equal(out["color"], ref), float(out["loss"]))
for v in leaves.values(): v.grad.zero_()
g.replay(); torch.cuda.synchronize()
print("grad relerr", float((leaves["means3D"].grad - gref).abs().max() / gref.abs().max()))
