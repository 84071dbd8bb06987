"""Does rendering two views concurrently on two CUDA streams raise throughput? (the binning stages are latency-bound and
leave most SMs idle; the blend kernels of another view could fill them)  Development aid; prints frames/s for 1..3 streams."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "cf-3dgs_b200")); sys.path.insert(0, ROOT)
import torch, gsr_synth as S, diff_gaussian_rasterization as D
dev = torch.device("cuda:0"); W, H, P = 1920, 1080, 1_000_000
cam = S.make_camera(W, H); sc = S.make_gvol(P, W, H, seed=5)
params = {k: sc[k].to(dev) for k in ["means3D", "opacities", "shs", "scales", "rotations"]}
gc = S.cotangents(W, H, seed=1, dense=False)["grad_color"].to(dev)
rs = D.GaussianRasterizationSettings(H, W, cam["tanfovx"], cam["tanfovy"], torch.zeros(3, device=dev), 1.0, cam["viewmatrix"].to(dev),
                                     cam["projmatrix"].to(dev), 3, cam["campos"].to(dev), False, False)
def step():
    lv = {k: v.detach().requires_grad_(True) for k, v in params.items()}
    m2d = torch.zeros_like(lv["means3D"], requires_grad=True)
    D.GaussianRasterizer(rs)(lv["means3D"], m2d, lv["opacities"], shs=lv["shs"], scales=lv["scales"], rotations=lv["rotations"])[0].backward(gc)
for sync_free in (False, True):
    D.set_sync_free(sync_free, slack=1.5)
    for n in (1, 2, 3):
        streams = [torch.cuda.Stream() for _ in range(n)]
        for s in streams:
            with torch.cuda.stream(s):
                for _ in range(3): step()
        torch.cuda.synchronize()
        K = 30
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for s in streams: s.wait_event(e0)
        for _ in range(K):
            for s in streams:
                with torch.cuda.stream(s): step()
        for s in streams: torch.cuda.current_stream().wait_stream(s)
        e1.record(); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1)
        print(f"sync_free={sync_free} streams={n}: {n * K / ms * 1e3:.1f} frames/s ({ms / (n * K):.3f} ms per frame)", flush=True)
print("status", D.sync_free_status())
