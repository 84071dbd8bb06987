#!/usr/bin/env python
"""CF-3DGS global optimisation steps on a synthetic Tanks-scale sequence (BASELINE.json configs[2]).

Shape of the reference's GLOBAL STEPS (trainer/cf3dgs_trainer.py:324-372): a global model (~1M Gaussians, the union of
per-frame pixel-aligned clouds) is optimised with Adam for `--iters` iterations, each against ONE randomly chosen
earlier frame rendered through that frame's (fixed) estimated camera; loss = 0.8*L1 + 0.2*(1-SSIM); SH degree grows
0 -> 3.  Densification / pruning (optimizer surgery, every 100 iterations in the reference) is outside the rasterizer
path and is not reproduced; the Gaussian count stays fixed.

  python examples/cf3dgs_global.py [--frames 8] [--points-per-frame 130000] [--iters 200] [--impl ours|reference]
prints one JSON line with global-step iterations/s.
"""
from __future__ import annotations

import argparse
import json
import math
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "cf-3dgs_b200")); sys.path.insert(0, ROOT); sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))

import numpy as np
import torch
import torch.nn.functional as F

import gsr_synth as S
import cf3dgs_pair as CP


def frame_pose(i, n):
    """Smooth camera trajectory: small yaw + sideways drift per frame (world-to-camera)."""
    a = math.radians(0.5 * i)
    R = np.array([[math.cos(a), 0, math.sin(a)], [0, 1, 0], [-math.sin(a), 0, math.cos(a)]])
    t = np.array([0.03 * i, 0.0, 0.01 * i])
    return R, t


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--frames", type=int, default=8)
    ap.add_argument("--points-per-frame", type=int, default=130_000)
    ap.add_argument("--iters", type=int, default=200)
    ap.add_argument("--width", type=int, default=960)
    ap.add_argument("--height", type=int, default=540)
    ap.add_argument("--sh-degree", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--loss", default=None, choices=["fused", "torch"])
    ap.add_argument("--profile", action="store_true", help="print the per-stage CUDA-event times of one iteration (ours)")
    ap.add_argument("--graph", action="store_true",
                    help="sync-free forward + one CUDA graph per training view holding the whole iteration (ours, fused model)")
    ap.add_argument("--model", default=None, choices=["fused", "torch"],
                    help="fused activation/Adam/statistics kernels (default for ours) or torch.optim.Adam + autograd")
    args = ap.parse_args()
    if args.model is None:
        args.model = "fused" if args.impl == "ours" else "torch"
    if args.loss is None:
        args.loss = "fused" if args.impl == "ours" else "torch"
    real_stdout = os.dup(1); os.dup2(2, 1)
    dev = torch.device("cuda", 0)
    W, H = args.width, args.height
    backend = CP.make_backend(args.impl, dev)
    loss_fn = CP.make_loss(args.loss)
    window = CP._gauss_window(device=dev)
    # ground-truth scene + per-frame cameras and target images
    gt = S.make_gpix(W, H, seed=3, target_P=250_000)
    gt_t = {k: gt[k].to(dev) for k in ["means3D", "shs", "scales", "rotations"]}
    gt_opa = torch.full((gt_t["means3D"].shape[0], 1), 0.95, device=dev)
    eye, zero6 = torch.eye(4, device=dev), torch.zeros(6, device=dev)
    cams, targets = [], []
    for i in range(args.frames):
        R, t = frame_pose(i, args.frames)
        cam = S.make_camera(W, H, R=R, t=t)
        rs = backend["settings"](cam, 0)
        with torch.no_grad():
            targets.append(backend["render"](rs, zero6, gt_t["means3D"], eye, gt_opa, shs=gt_t["shs"], scales=gt_t["scales"],
                                             rotations=gt_t["rotations"])[0].clamp(0, 1))
        cams.append(cam)
    # global model: union of per-frame pixel-aligned clouds, moved into the world frame of camera 0
    xs, ss, cs = [], [], []
    for i in range(args.frames):
        c = S.make_gpix(W, H, seed=3, target_P=args.points_per_frame)
        R, t = frame_pose(i, args.frames)
        Rt = torch.tensor(R, dtype=torch.float32); tt = torch.tensor(t, dtype=torch.float32)
        xs.append((c["means3D"] - tt) @ Rt)           # camera i -> world: R^T (p - t)
        ss.append(c["scales"]); cs.append(c["shs"])
    xyz = torch.cat(xs).to(dev).requires_grad_(True)
    P = xyz.shape[0]
    shs0 = torch.cat(cs).to(dev)
    f_dc = shs0[:, :1].clone().requires_grad_(True)
    f_rest = shs0[:, 1:].clone().requires_grad_(True)
    log_s = torch.cat(ss).to(dev).log().requires_grad_(True)
    rot = torch.zeros(P, 4, device=dev); rot[:, 0] = 1; rot.requires_grad_(True)
    opa_raw = torch.full((P, 1), math.log(0.1 / 0.9), device=dev).requires_grad_(True)
    opt = torch.optim.Adam([{"params": [xyz], "lr": 0.00016}, {"params": [f_dc], "lr": 0.0025}, {"params": [f_rest], "lr": 0.000125},
                            {"params": [opa_raw], "lr": 0.05}, {"params": [log_s], "lr": 0.005}, {"params": [rot], "lr": 0.001}],
                           eps=1e-15, fused=True)
    settings = [backend["settings"](c, args.sh_degree) for c in cams]
    order = np.random.RandomState(0).randint(0, args.frames, size=args.iters + 10)

    fused_model = None
    if args.model == "fused":
        import gsr_model
        fused_model = gsr_model.FusedGaussianModel(
            xyz, torch.cat([f_dc, f_rest], 1), opa_raw, log_s, rot,
            dict(xyz=0.00016, f_dc=0.0025, f_rest=0.000125, opacity=0.05, scaling=0.005, rotation=0.001),
            device_step=bool(args.graph))

    def fused_iteration(k):
        return fused_iteration_frame(int(order[k]))

    def fused_iteration_frame(f):
        m = fused_model
        act = m.activated()
        x = m.xyz.detach().requires_grad_(True)
        sh = m.features.detach().requires_grad_(True)
        o, sc, r = act["opacities"].requires_grad_(True), act["scales"].requires_grad_(True), act["rotations"].requires_grad_(True)
        m2d = torch.zeros_like(x, requires_grad=True)
        import diff_gaussian_rasterization as D
        img, radii, _, _ = D.GaussianRasterizer(settings[f])(x, m2d, o, shs=sh, scales=sc, rotations=r)
        loss = loss_fn(img, targets[f], window)
        loss.backward()
        m.step(x.grad, sh.grad, o.grad, sc.grad, r.grad, d_means2D=m2d.grad, radii=radii)
        return loss

    def iteration(k):
        if fused_model is not None:
            return fused_iteration(k)
        f = int(order[k])
        img = backend["render"](settings[f], zero6, xyz, eye, torch.sigmoid(opa_raw), shs=torch.cat([f_dc, f_rest], 1),
                                scales=torch.exp(log_s), rotations=F.normalize(rot))[0]
        loss = loss_fn(img, targets[f], window)
        loss.backward()
        opt.step(); opt.zero_grad(set_to_none=True)
        return loss

    if args.profile and args.impl == "ours":
        import ctypes
        import diff_gaussian_rasterization as D
        for k in range(3):
            iteration(k)
        n_st = D.lib.gsr_profile_count()
        acc = [0.0] * n_st
        D.lib.gsr_profile_enable(1)
        for k in range(3, 3 + 8):
            iteration(k)
            buf = (ctypes.c_float * n_st)(); D.lib.gsr_profile_read(buf, n_st)
            acc = [a + max(b, 0.0) for a, b in zip(acc, buf[:])]
        D.lib.gsr_profile_enable(0)
        print({D.lib.gsr_profile_name(i).decode(): round(acc[i] / 8, 4) for i in range(n_st)}, file=sys.stderr)
    overflow = None
    if args.graph:
        if fused_model is None:
            raise SystemExit("--graph needs --impl ours --model fused")
        import diff_gaussian_rasterization as D
        D.set_sync_free(True, slack=2.0)
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            for f in range(args.frames):     # learns the binning capacity over all views (raised if a view needs more)
                first = fused_iteration_frame(f)
            for k in range(10):
                first = iteration(k)
        torch.cuda.current_stream().wait_stream(side)
        D.sync_free_status()
        graphs = []
        for f in range(args.frames):
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g):
                lf = fused_iteration_frame(f)
            graphs.append((g, lf))
        torch.cuda.synchronize(dev)
        t0 = time.time()
        for k in range(10, 10 + args.iters):
            g, loss = graphs[int(order[k])]
            g.replay()
        torch.cuda.synchronize(dev)
        dt = time.time() - t0
        st = D.sync_free_status()
        overflow = bool(st[1]) if st else None
    else:
        for k in range(10):
            first = iteration(k)
        torch.cuda.synchronize(dev)
        t0 = time.time()
        for k in range(10, 10 + args.iters):
            loss = iteration(k)
        torch.cuda.synchronize(dev)
        dt = time.time() - t0
    line = {"workload": "cf3dgs_global_steps", "impl": args.impl, "loss": args.loss, "model": args.model, "P": P, "frames": args.frames, "W": W, "H": H,
            "sh_degree": args.sh_degree, "iters": args.iters, "iters_per_s": args.iters / dt, "ms_per_iter": dt / args.iters * 1e3,
            "first_loss": float(first.detach()), "last_loss": float(loss.detach()), "graph": bool(args.graph)}
    if overflow is not None:
        line["binning_overflow"] = overflow
    os.write(real_stdout, (json.dumps(line) + "\n").encode())


if __name__ == "__main__":
    main()
