"""Sort / raster roofline sweep against the Gaussian count (BASELINE.json configs[4]: up to the 5M stress render at 1080p).

For every P: fwd+bwd of the G-vol scene at 1920x1080, SH degree 3, colour cotangent only -- per-stage CUDA-event times
(gsr_profile_*), the work counters, achieved GB/s (HBM-bound stages) or TFLOP/s (blend kernels), and, when oracle/_ref
travelled to the box, the reference build's fwd+bwd time on the same inputs.  One JSON line per P on stdout.

  python tools/sweep.py [P ...]
"""
import ctypes
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "cf-3dgs_b200")); sys.path.insert(0, ROOT)
import numpy as np
import torch
import gsr_synth as S
import diff_gaussian_rasterization as D

W, H = 1920, 1080
dev = torch.device("cuda:0")
lib = D.lib
sizes = [int(a) for a in sys.argv[1:]] or [100_000, 300_000, 1_000_000, 2_000_000, 3_000_000, 5_000_000]
peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))) if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else {}
hbm_peak = float(peaks.get("hbm_gbs", 6650.0))
cam = S.make_camera(W, H)
gc = S.cotangents(W, H, seed=1, dense=False)["grad_color"].to(dev)
bg = torch.zeros(3, device=dev)
rs = D.GaussianRasterizationSettings(image_height=H, image_width=W, tanfovx=cam["tanfovx"], tanfovy=cam["tanfovy"], bg=bg,
                                     scale_modifier=1.0, viewmatrix=cam["viewmatrix"].to(dev), projmatrix=cam["projmatrix"].to(dev),
                                     sh_degree=3, campos=cam["campos"].to(dev), prefiltered=False, debug=False)
try:
    from oracle import ref_cuda as RC
    have_ref = RC.available()
except Exception:
    have_ref = False


def timed(fn, n=10, warm=3):
    for _ in range(warm):
        fn()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); e0.record()
    for _ in range(n):
        fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


for P in sizes:
    sc = S.make_gvol(P, W, H, seed=5)
    params = {k: sc[k].to(dev) for k in ["means3D", "opacities", "shs", "scales", "rotations"]}
    holder = {}

    def step():
        lv = {k: v.detach().requires_grad_(True) for k, v in params.items()}
        m2d = torch.zeros_like(lv["means3D"], requires_grad=True)
        out = D.GaussianRasterizer(rs)(lv["means3D"], m2d, lv["opacities"], shs=lv["shs"], scales=lv["scales"], rotations=lv["rotations"])
        out[0].backward(gc)
        holder["out"] = out

    ms = timed(step)
    n_st = lib.gsr_profile_count()
    names = [lib.gsr_profile_name(i).decode() for i in range(n_st)]
    acc = np.zeros(n_st)
    lib.gsr_profile_enable(1)
    for _ in range(5):
        step()
        buf = (ctypes.c_float * n_st)(); lib.gsr_profile_read(buf, n_st)
        acc += np.maximum(np.array(buf[:]), 0.0)
    lib.gsr_profile_enable(0)
    st = {n: float(a / 5) for n, a in zip(names, acc)}
    # work counters from one more forward (the saved workspaces are released by backward)
    lv = {k: v.detach().requires_grad_(True) for k, v in params.items()}
    holder["out"] = D.GaussianRasterizer(rs)(lv["means3D"], torch.zeros_like(lv["means3D"], requires_grad=True), lv["opacities"],
                                             shs=lv["shs"], scales=lv["scales"], rotations=lv["rotations"])
    fn = holder["out"][0].grad_fn
    R = fn.num_rendered
    counts = torch.zeros(4, dtype=torch.int64, device=dev)
    lib.gsr_debug_pair_counts(P, W, H, R, fn.saved_tensors[7].data_ptr(), fn.saved_tensors[8].data_ptr(), fn.saved_tensors[9].data_ptr(),
                              counts.data_ptr(), torch.cuda.current_stream().cuda_stream)
    in_range, blended, _, _ = [int(x) for x in counts.tolist()]
    vis = int((holder["out"][1] > 0).sum())
    tiles = ((W + 15) // 16) * ((H + 15) // 16)
    sort_ms = st["depth_sort"] + st["offsets_scan"] + st["emit"] + st["tile_sort"] + st.get("tile_ranges", 0.0)
    sort_bytes = P * 16 + P * 8 + (P * 20 + R * 12) + R * 24 + R * 8 + tiles * 8          # SURVEY.md 8d floors
    line = {"P": P, "visible": vis, "num_rendered": R, "pairs_blended": blended, "ms_fwd_bwd": ms, "frames_per_s": 1000.0 / ms,
            "gaussians_per_s": P * 1000.0 / ms, "stages_ms": {k: round(v, 4) for k, v in st.items()},
            "preprocess_gbs": (vis * 311 + (P - vis) * 20) / 1e9 / (st["preprocess"] / 1e3),
            "preprocess_backward_gbs": vis * 627 / 1e9 / (st["preprocess_backward"] / 1e3),
            "binning_ms": sort_ms, "binning_gbs": sort_bytes / 1e9 / (sort_ms / 1e3), "hbm_peak_gbs": hbm_peak,
            "render_forward_tflops": (in_range * 24 + blended * 16) / 1e12 / (st["render_forward"] / 1e3),
            "render_backward_tflops": (blended * 75 + (in_range - blended) * 24) / 1e12 / (st["render_backward"] / 1e3)}
    if have_ref:
        def ref_step():
            r = RC.forward(params["means3D"], params["opacities"], shs=params["shs"], scales=params["scales"], rotations=params["rotations"],
                           viewmatrix=rs.viewmatrix, projmatrix=rs.projmatrix, campos=rs.campos, bg=bg, W=W, H=H,
                           tanfovx=cam["tanfovx"], tanfovy=cam["tanfovy"], sh_degree=3)
            RC.backward(r, gc, torch.zeros(1, H, W, device=dev), torch.zeros(1, H, W, device=dev))
        line["reference_ms_fwd_bwd"] = timed(ref_step, n=5, warm=2)
        line["speedup_vs_reference"] = line["reference_ms_fwd_bwd"] / ms
    print(json.dumps(line), flush=True)
    del params, sc, holder
    torch.cuda.empty_cache()
