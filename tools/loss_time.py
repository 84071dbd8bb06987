"""Time of the fused loss kernels at 960x540 and 1920x1080 (development aid)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "cf-3dgs_b200")); sys.path.insert(0, ROOT)
import torch, diff_gaussian_rasterization as D
lib = D.lib
for H, W in [(540, 960), (1080, 1920)]:
    img = torch.rand(3, H, W, device="cuda"); gt = torch.rand(3, H, W, device="cuda")
    scratch = torch.empty(lib.gsr_photo_loss_scratch_bytes(3, H, W), dtype=torch.uint8, device="cuda")
    out = torch.zeros(3, device="cuda"); one = torch.ones(1, device="cuda"); dimg = torch.empty_like(img)
    s = torch.cuda.current_stream().cuda_stream
    def fwd(): lib.gsr_photo_loss_forward(3, H, W, img.data_ptr(), gt.data_ptr(), 0.2, 1, scratch.data_ptr(), out.data_ptr(), s)
    def bwd(): lib.gsr_photo_loss_backward(3, H, W, img.data_ptr(), gt.data_ptr(), 0.2, 1, scratch.data_ptr(), one.data_ptr(), dimg.data_ptr(), s)
    for name, fn in [("forward", fwd), ("backward", bwd)]:
        for _ in range(10): fn()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize(); e0.record()
        for _ in range(200): fn()
        e1.record(); torch.cuda.synchronize()
        print(f"{W}x{H} {name}: {e0.elapsed_time(e1)/200*1000:.1f} us  loss {float(out[0]):.6f}")
