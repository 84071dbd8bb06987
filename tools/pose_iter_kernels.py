"""Kernel list of one CF-3DGS pose-fit iteration of gsr_pose.PoseOptimizer (development aid): which launches are the
library's and whether any PyTorch op is left around them.  (Kernel durations reported by a profiler include the time a
programmatically launched kernel spends waiting for its predecessor: read the COUNTS here, and the stage timers of
gsr_profile_* for times.)"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "cf-3dgs_b200")); sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "examples"))
import torch, gsr_synth as S, diff_gaussian_rasterization as D, gsr_pose as GP
import cf3dgs_pair as CP
from torch.profiler import profile, ProfilerActivity
dev = torch.device("cuda:0"); W, H = 960, 540
cam = S.make_camera(W, H); sc = S.make_gpix(W, H, seed=2, target_P=250000)
backend = CP.make_backend("ours", dev); rs = backend["settings"](cam, 0)
t = {k: sc[k].to(dev) for k in ["means3D", "shs", "scales", "rotations"]}
opa = torch.full((t["means3D"].shape[0], 1), 0.5, device=dev)
loss_fn = CP.make_loss("fused"); gt = torch.rand(3, H, W, device=dev)
D.set_sync_free(True, slack=1.5)
popt = GP.PoseOptimizer(rs, t["means3D"], opa, shs=t["shs"], scales=t["scales"], rotations=t["rotations"], lr=1e-3)
def iteration():
    popt.step(gt)
for _ in range(5): iteration()
torch.cuda.synchronize()
with profile(activities=[ProfilerActivity.CUDA]) as prof:
    for _ in range(10): iteration()
    torch.cuda.synchronize()
rows = [(e.key, e.count, e.device_time_total) for e in prof.key_averages()]
ours = sum(r[2] for r in rows if "gsr::" in r[0]); other = sum(r[2] for r in rows if "gsr::" not in r[0])
n_ours = sum(r[1] for r in rows if "gsr::" in r[0]); n_other = sum(r[1] for r in rows if "gsr::" not in r[0])
print(f"per iteration: library kernels {n_ours/10:.0f} launches {ours/10:.0f} us | other (PyTorch ops, memsets, copies) {n_other/10:.0f} launches {other/10:.0f} us")
for k, c, tt in sorted(rows, key=lambda r: -r[2])[:40]:
    print(f"{tt/10:8.1f} us {c/10:5.1f} x  {k[:110]}")
