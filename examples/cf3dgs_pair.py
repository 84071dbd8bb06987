#!/usr/bin/env python
"""CF-3DGS local frame-pair optimisation on synthetic "Tanks-like" frames (BASELINE.json configs[1] and [3]).

Reproduces the tensor traffic of one `add_view_v2` step of the reference trainer
(trainer/cf3dgs_trainer.py:219-297) without its un-installable dependencies:
  1. LOCAL GAUSSIAN FIT  (:225-259): a fresh local model (one Gaussian per k-th pixel of frame i-1, opacity 0.1,
     SH degree 0 with M=16) is fitted to frame i-1 with Adam, identity camera,
     loss = 0.8*L1 + 0.2*(1-SSIM) (trainer/losses.py:98-136);
  2. LOCAL POSE FIT      (:262-285): Gaussians frozen, only the rigid pose is optimised against frame i for 300
     iterations; the pose gradient comes fused out of the rasterizer backward (gsr_pose.rasterize_with_pose).
Frame i is rendered from the ground-truth scene under a known small SE(3) offset, so the fit has a known answer.

Multi-GPU: frame pairs are independent -> pair p runs on rank p mod world (no data-path collective); at the end the
4x4 relative poses are gathered and chained (gsr_parallel).  Launch with torchrun for N>1.

  python examples/cf3dgs_pair.py [--pairs 1] [--fit-iters 200] [--pose-iters 300] [--impl ours|reference]
prints one JSON line with iters/s of both phases and the pose error.
"""
from __future__ import annotations

import argparse
import json
import math
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "cf-3dgs_b200"))
sys.path.insert(0, ROOT)

import numpy as np
import torch
import torch.nn.functional as F

import gsr_synth as S


def _gauss_window(size=11, sigma=1.5, device="cpu"):
    g = torch.tensor([math.exp(-(x - size // 2) ** 2 / (2 * sigma ** 2)) for x in range(size)], device=device)
    g = (g / g.sum()).unsqueeze(1)
    w2 = (g @ g.t()).unsqueeze(0).unsqueeze(0)
    return w2.expand(3, 1, size, size).contiguous()


def ssim(img1, img2, window):
    """utils/loss_utils.py ssim(): 11x11 Gaussian window, sigma 1.5."""
    mu1 = F.conv2d(img1, window, padding=5, groups=3)
    mu2 = F.conv2d(img2, window, padding=5, groups=3)
    mu1_sq, mu2_sq, mu12 = mu1 * mu1, mu2 * mu2, mu1 * mu2
    s1 = F.conv2d(img1 * img1, window, padding=5, groups=3) - mu1_sq
    s2 = F.conv2d(img2 * img2, window, padding=5, groups=3) - mu2_sq
    s12 = F.conv2d(img1 * img2, window, padding=5, groups=3) - mu12
    C1, C2 = 0.01 ** 2, 0.03 ** 2
    return (((2 * mu12 + C1) * (2 * s12 + C2)) / ((mu1_sq + mu2_sq + C1) * (s1 + s2 + C2))).mean()


def photometric_loss(img, gt, window, lam=0.2):
    """PyTorch formulation (what the reference trainer runs): expects the clamped image."""
    img = img.clamp(0, 1)
    return (1.0 - lam) * (img - gt).abs().mean() + lam * (1.0 - ssim(img.unsqueeze(0), gt.unsqueeze(0), window))


def make_loss(kind):
    if kind == "fused":
        import gsr_loss
        return lambda img, gt, window: gsr_loss.photometric_loss(img, gt, 0.2, clamp=True)
    return photometric_loss


def true_offset(p):
    a = math.radians(0.6 + 0.2 * (p % 5))
    return torch.tensor([[math.cos(a), 0, math.sin(a), 0.02 + 0.002 * (p % 3)], [0, 1, 0, -0.01],
                         [-math.sin(a), 0, math.cos(a), 0.015], [0, 0, 0, 1]], dtype=torch.float32)


def run_pair(p, args, dev, backend):
    import gsr_pose as GP
    loss_fn = make_loss(args.loss)
    W, H = args.width, args.height
    cam = S.make_camera(W, H)
    sc = S.make_gpix(W, H, seed=2 + p, target_P=args.points)
    rs = backend["settings"](cam, 0)
    P = sc["means3D"].shape[0]
    window = _gauss_window(device=dev)
    xyz_gt = sc["means3D"].to(dev)
    eye = torch.eye(4, device=dev)
    zero6 = torch.zeros(6, device=dev)
    gt_kw = dict(shs=sc["shs"].to(dev), scales=sc["scales"].to(dev), rotations=sc["rotations"].to(dev))
    opa_gt = torch.full((P, 1), 0.95, device=dev)
    with torch.no_grad():
        frame0 = backend["render"](rs, zero6, xyz_gt, eye, opa_gt, **gt_kw)[0].clamp(0, 1)
        T_true = true_offset(p).to(dev)
        frame1 = backend["render"](rs, zero6, xyz_gt, T_true, opa_gt, **gt_kw)[0].clamp(0, 1)

    # ---- 1. local Gaussian fit (raw parameters + activations as in scene/gaussian_model_cf.py:41-56) ----
    xyz = xyz_gt.clone().requires_grad_(True)
    f_dc = torch.zeros(P, 1, 3, device=dev).requires_grad_(True)
    f_rest = torch.zeros(P, 15, 3, device=dev).requires_grad_(True)
    log_s = sc["scales"].to(dev).log().requires_grad_(True)
    rot = sc["rotations"].to(dev).clone().requires_grad_(True)
    opa_raw = torch.full((P, 1), math.log(0.1 / 0.9), device=dev).requires_grad_(True)   # inverse_sigmoid(0.1)
    groups = [{"params": [xyz], "lr": 0.0}, {"params": [f_dc], "lr": 0.0025}, {"params": [f_rest], "lr": 0.000125},
              {"params": [opa_raw], "lr": 0.05}, {"params": [log_s], "lr": 0.005}, {"params": [rot], "lr": 0.001}]

    def fit_iteration():
        img = backend["render"](rs, zero6, xyz, eye, torch.sigmoid(opa_raw), shs=torch.cat([f_dc, f_rest], 1),
                                scales=torch.exp(log_s), rotations=F.normalize(rot))[0]
        loss = loss_fn(img, frame0, window)
        loss.backward()
        return loss

    fused_model = None
    if args.model == "fused":
        # activations + their chain rule + Adam in the library's fused model step (gsr_model.py) instead of PyTorch ops
        import gsr_model
        import diff_gaussian_rasterization as D
        fused_model = gsr_model.FusedGaussianModel(
            xyz, torch.cat([f_dc, f_rest], 1), opa_raw, log_s, rot,
            dict(xyz=0.0, f_dc=0.0025, f_rest=0.000125, opacity=0.05, scaling=0.005, rotation=0.001),
            device_step=True)

        def fit_iteration_fused():
            mdl = fused_model
            act = mdl.activated()
            x = mdl.xyz.detach().requires_grad_(True)
            sh = mdl.shs(0).detach().requires_grad_(True)     # local models stay at SH degree 0: the dense coefficient-0 plane
            o, sc_, r = act["opacities"].requires_grad_(True), act["scales"].requires_grad_(True), act["rotations"].requires_grad_(True)
            m2d = torch.zeros_like(x, requires_grad=True)
            img = D.GaussianRasterizer(rs)(x, m2d, o, shs=sh, scales=sc_, rotations=r)[0]
            loss = loss_fn(img, frame0, window)
            loss.backward()
            mdl.step(x.grad, sh.grad, o.grad, sc_.grad, r.grad, sh_degree=0)
            return loss

    if fused_model is not None and args.graph:
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            for _ in range(3):
                fit_iteration_fused()
        torch.cuda.current_stream().wait_stream(side)
        fit_graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(fit_graph):
            loss = fit_iteration_fused()
        torch.cuda.synchronize(dev)
        t0 = time.time()
        for it in range(args.fit_iters):
            fit_graph.replay()
    elif fused_model is not None:
        torch.cuda.synchronize(dev)
        t0 = time.time()
        for it in range(args.fit_iters):
            loss = fit_iteration_fused()
    elif args.graph:
        # one CUDA graph per fit iteration: render, loss, backward, fused multi-tensor Adam
        opt = torch.optim.Adam(groups, eps=1e-15, capturable=True, fused=True)
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            for _ in range(3):
                fit_iteration(); opt.step(); opt.zero_grad(set_to_none=True)
        torch.cuda.current_stream().wait_stream(side)
        fit_graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(fit_graph):
            loss = fit_iteration()
            opt.step()
            opt.zero_grad(set_to_none=False)
        torch.cuda.synchronize(dev)
        t0 = time.time()
        for it in range(args.fit_iters):
            fit_graph.replay()
    else:
        opt = torch.optim.Adam(groups, eps=1e-15)
        torch.cuda.synchronize(dev)
        t0 = time.time()
        for it in range(args.fit_iters):
            loss = fit_iteration()
            opt.step(); opt.zero_grad(set_to_none=True)
    torch.cuda.synchronize(dev)
    fit_s = time.time() - t0
    fit_loss = float(loss.detach())

    if fused_model is not None:   # hand the fitted raw parameters to the pose stage below
        f_dc, f_rest = fused_model.features[:, :1], fused_model.features[:, 1:]
        opa_raw, log_s, rot, xyz = fused_model.opacity_raw, fused_model.log_scales, fused_model.rot_raw, fused_model.xyz

    # ---- 2. local pose fit: Gaussians frozen, pose only ----
    with torch.no_grad():
        fr = dict(shs=torch.cat([f_dc, f_rest], 1).detach(), scales=torch.exp(log_s).detach(),
                  rotations=F.normalize(rot).detach())
        opa = torch.sigmoid(opa_raw).detach()
        xyz_f = xyz.detach()
    T = torch.eye(4, device=dev)
    m = torch.zeros(6, device=dev); v = torch.zeros(6, device=dev)
    lr, b1, b2 = args.pose_lr, 0.9, 0.999
    fused_pose = args.impl == "ours" and args.loss == "fused" and getattr(args, "pose_loop", "fused") == "fused"
    if fused_pose:
        # the library's pose loop: rigid render, fused loss, twist-only backward, Adam + SE(3) retraction on the device --
        # no PyTorch kernel in the iteration (gsr_pose.PoseOptimizer); as one replayed CUDA graph with --graph
        popt = GP.PoseOptimizer(rs, xyz_f, opa, shs=fr["shs"], scales=fr["scales"], rotations=fr["rotations"], lr=lr,
                                betas=(b1, b2), eps=1e-8)
        if args.graph:
            side = torch.cuda.Stream()
            side.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(side):
                for _ in range(3):
                    popt.step(frame1)
            torch.cuda.current_stream().wait_stream(side)
            popt.reset()
            graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(graph):
                loss = popt.step(frame1)
            popt.reset()
            torch.cuda.synchronize(dev)
            t0 = time.time()
            for it in range(args.pose_iters):
                graph.replay()
        else:
            torch.cuda.synchronize(dev)
            t0 = time.time()
            for it in range(args.pose_iters):
                loss = popt.step(frame1)
        torch.cuda.synchronize(dev)
        pose_s = time.time() - t0
        T = popt.T
    elif args.graph:
        # whole pose iteration (render, loss, backward, Adam on the twist, pose retraction) as ONE CUDA graph:
        # needs the sync-free forward (no host read of num_rendered inside the loop)
        xi = torch.zeros(6, device=dev, requires_grad=True)
        t_dev = torch.zeros((), device=dev)
        b1_dev, b2_dev = torch.full((), b1, device=dev), torch.full((), b2, device=dev)

        def iteration():
            img = backend["render"](rs, xi, xyz_f, T, opa, **fr)[0]
            loss = loss_fn(img, frame1, window)
            loss.backward()
            g = xi.grad
            t_dev.add_(1.0)
            m.mul_(b1).add_(g, alpha=1 - b1)
            v.mul_(b2).addcmul_(g, g, value=1 - b2)
            step = -lr * (m / (1 - torch.pow(b1_dev, t_dev))) / ((v / (1 - torch.pow(b2_dev, t_dev))).sqrt() + 1e-8)
            T.copy_(GP.se3_exp_closed(step) @ T)
            g.zero_()
            return loss

        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            for _ in range(3):
                iteration()
        torch.cuda.current_stream().wait_stream(side)
        T.copy_(torch.eye(4, device=dev)); m.zero_(); v.zero_(); t_dev.zero_(); xi.grad = None
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph):
            loss = iteration()
        T.copy_(torch.eye(4, device=dev)); m.zero_(); v.zero_(); t_dev.zero_()
        torch.cuda.synchronize(dev)
        t0 = time.time()
        for it in range(args.pose_iters):
            graph.replay()
        torch.cuda.synchronize(dev)
        pose_s = time.time() - t0
    else:
        torch.cuda.synchronize(dev)
        t0 = time.time()
        for it in range(args.pose_iters):
            xi = torch.zeros(6, device=dev, requires_grad=True)
            img = backend["render"](rs, xi, xyz_f, T, opa, **fr)[0]
            loss = loss_fn(img, frame1, window)
            loss.backward()
            g = xi.grad
            m = b1 * m + (1 - b1) * g; v = b2 * v + (1 - b2) * g * g
            step = -lr * (m / (1 - b1 ** (it + 1))) / ((v / (1 - b2 ** (it + 1))).sqrt() + 1e-8)
            T = (GP.se3_exp(step) @ T)
        torch.cuda.synchronize(dev)
        pose_s = time.time() - t0
    err_t = float((T[:3, 3] - T_true[:3, 3]).norm())
    err_r = float(torch.acos(((T[:3, :3].t() @ T_true[:3, :3]).trace().clamp(-1, 3) - 1) / 2).abs()) * 180 / math.pi
    return dict(pair=p, P=P, fit_s=fit_s, pose_s=pose_s, fit_loss=fit_loss, pose_loss=float(loss.detach()),
                err_t=err_t, err_r_deg=err_r, T=T)


def make_backend(impl, dev):
    import gsr_pose as GP
    if impl == "ours":
        import diff_gaussian_rasterization as D

        def settings(cam, deg):
            return D.GaussianRasterizationSettings(cam["H"], cam["W"], cam["tanfovx"], cam["tanfovy"], torch.zeros(3, device=dev),
                                                   1.0, cam["viewmatrix"].to(dev), cam["projmatrix"].to(dev), deg,
                                                   cam["campos"].to(dev), False, False)

        def render(rs, xi, xyz, T, opa, shs, scales, rotations):
            return GP.rasterize_with_pose(rs, xi, xyz, T, opa, shs=shs, scales=scales, rotations=rotations)
        return dict(settings=settings, render=render)
    from oracle import ref_cuda as RC

    class _Ref(torch.autograd.Function):
        @staticmethod
        def forward(ctx, means3D, opa, shs, scales, rots, rs):
            st = RC.forward(means3D, opa, shs=shs, scales=scales, rotations=rots, **rs)
            # the state goes through save_for_backward, as the reference's own wrapper does: keeping the output tensors on
            # ctx directly would tie node and outputs into a reference cycle that only the garbage collector breaks, and a
            # GB of workspaces per iteration then piles up between collections
            ctx.save_for_backward(st["radii"], st["geom"], st["binning"], st["img"], st["alpha"])
            ctx.meta = (st["num_rendered"], st["_args"])
            return st["color"], st["depth"], st["alpha"]

        @staticmethod
        def backward(ctx, gc, gd, ga):
            radii, geom, binning, img, alpha = ctx.saved_tensors
            st = dict(radii=radii, geom=geom, binning=binning, img=img, alpha=alpha, num_rendered=ctx.meta[0], _args=ctx.meta[1])
            z = torch.zeros_like(alpha)
            g = RC.backward(st, gc.contiguous(), z if gd is None else gd.contiguous(), z if ga is None else ga.contiguous())
            return g["means3D"], g["opacities"], g["sh"], g["scales"], g["rotations"], None

    def settings(cam, deg):
        return dict(viewmatrix=cam["viewmatrix"].to(dev), projmatrix=cam["projmatrix"].to(dev), campos=cam["campos"].to(dev),
                    bg=torch.zeros(3, device=dev), W=cam["W"], H=cam["H"], tanfovx=cam["tanfovx"], tanfovy=cam["tanfovy"],
                    sh_degree=deg)

    def render(rs, xi, xyz, T, opa, shs, scales, rotations):
        # the reference obtains the pose gradient by autograd through the point transform (lietorch there, matrix_exp here)
        M = GP.se3_exp(xi) @ T
        means = (xyz @ M[:3, :3].t() + M[:3, 3]).contiguous()
        c, d, a = _Ref.apply(means, opa.contiguous(), shs.contiguous(), scales.contiguous(), rotations.contiguous(), rs)
        return c, None, d, a
    return dict(settings=settings, render=render)


def default_args(**kw):
    d = dict(pairs=1, fit_iters=200, pose_iters=300, points=250_000, width=960, height=540, pose_lr=1e-3, impl="ours",
             graph=False, sync_free=False, model=None, loss=None, pose_loop="fused")
    d.update(kw)
    if d["loss"] is None:
        d["loss"] = "fused" if d["impl"] == "ours" else "torch"
    if d["model"] is None:
        d["model"] = "fused" if d["impl"] == "ours" else "torch"
    return argparse.Namespace(**d)


def run_sharded(args, dev, rank=0, world=1):
    """Frame pairs sharded over the ranks (pair p on rank p mod world, no data-path collective), then one gather of
    the relative poses.  Returns (result dict on rank 0 / None elsewhere, binning overflow flag)."""
    import gsr_parallel as GPar
    backend = make_backend(args.impl, dev)
    if args.impl == "ours" and (args.graph or args.sync_free):
        import diff_gaussian_rasterization as D
        D.set_sync_free(True, slack=1.5)
    run_pair(0, argparse.Namespace(**{**vars(args), "fit_iters": 5, "pose_iters": 5}), dev, backend)   # warm-up
    mine = GPar.shard_pairs(args.pairs, rank, world)
    if world > 1:
        import torch.distributed as dist
        dist.barrier()
    torch.cuda.synchronize(dev)
    t0 = time.time()
    res = [run_pair(p, args, dev, backend) for p in mine]
    torch.cuda.synchronize(dev)
    local_s = time.time() - t0
    overflow = None
    if args.impl == "ours" and (args.graph or args.sync_free):
        # every sync-free render of the run (eager, captured, replayed) has OR-ed into the device's status buffer
        st = D.sync_free_status()
        overflow = bool(st[1]) if st else False
    rel = GPar.gather_relative_poses({r["pair"]: r["T"] for r in res}, args.pairs) if world > 1 else [r["T"] for r in res]
    poses = GPar.compose_poses(rel)
    if world > 1:
        import torch.distributed as dist
        t = torch.tensor([local_s, sum(r["fit_s"] for r in res), sum(r["pose_s"] for r in res),
                          max([r["err_t"] for r in res] + [0.0]), max([r["err_r_deg"] for r in res] + [0.0])], device=dev)
        tmax = t.clone(); dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
        wall, fit_s, pose_s, err_t, err_r = [float(x) for x in tmax.tolist()]
    else:
        wall, fit_s, pose_s = local_s, sum(r["fit_s"] for r in res), sum(r["pose_s"] for r in res)
        err_t, err_r = max(r["err_t"] for r in res), max(r["err_r_deg"] for r in res)
    if rank == 0:
        n_local = len(GPar.shard_pairs(args.pairs, 0, world))
        line = {"workload": "cf3dgs_local_pair", "impl": args.impl, "loss": args.loss, "model": args.model, "graph": bool(args.graph), "sync_free": bool(args.graph or args.sync_free),
                "n_gpus": world, "pairs": args.pairs, "P": res[0]["P"] if res else 0,
                "W": args.width, "H": args.height, "fit_iters": args.fit_iters, "pose_iters": args.pose_iters,
                "wall_s": wall, "pairs_per_s": args.pairs / wall,
                "pose_opt_iters_per_s": args.pairs * args.pose_iters / (pose_s * world) * world if pose_s > 0 else None,
                "pose_opt_iters_per_s_aggregate": args.pairs * args.pose_iters / pose_s if pose_s > 0 else None,
                "gaussian_fit_iters_per_s_aggregate": args.pairs * args.fit_iters / fit_s if fit_s > 0 else None,
                "max_pose_err_t": err_t, "max_pose_err_r_deg": err_r, "n_poses": len(poses),
                "fit_loss": res[0]["fit_loss"] if res else None, "pose_loss": res[0]["pose_loss"] if res else None}
        if overflow is not None:
            line["binning_overflow"] = overflow
        return line, overflow
    return None, overflow



def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--pairs", type=int, default=1)
    ap.add_argument("--fit-iters", type=int, default=200)
    ap.add_argument("--pose-iters", type=int, default=300)
    ap.add_argument("--points", type=int, default=250_000)
    ap.add_argument("--width", type=int, default=960)
    ap.add_argument("--height", type=int, default=540)
    ap.add_argument("--pose-lr", type=float, default=1e-3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--graph", action="store_true",
                    help="sync-free forward + the pose iteration captured as one CUDA graph (ours only)")
    ap.add_argument("--sync-free", action="store_true", help="sync-free forward without graph capture (ours only)")
    ap.add_argument("--pose-loop", default="fused", choices=["fused", "torch"],
                    help="pose fit: the library's loop (gsr_pose.PoseOptimizer, default for ours with the fused loss) or "
                         "autograd + PyTorch Adam / retraction ops")
    ap.add_argument("--model", default=None, choices=["fused", "torch"],
                    help="Gaussian fit: the library's fused model step (default for ours) or PyTorch activations + Adam")
    ap.add_argument("--loss", default=None, choices=["fused", "torch"],
                    help="fused CUDA L1+SSIM (default for --impl ours) or the PyTorch formulation (default for reference)")
    args = ap.parse_args()
    if args.loss is None:
        args.loss = "fused" if args.impl == "ours" else "torch"
    if args.model is None:
        args.model = "fused" if args.impl == "ours" else "torch"
    if args.model == "fused" and args.impl != "ours":
        raise SystemExit("--model fused needs --impl ours")
    world = int(os.environ.get("WORLD_SIZE", "1")); rank = int(os.environ.get("RANK", "0")); local = int(os.environ.get("LOCAL_RANK", "0"))
    real_stdout = os.dup(1); os.dup2(2, 1)
    if world > 1:
        import datetime
        import torch.distributed as dist
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1"); os.environ.setdefault("MASTER_PORT", "29500")
        dist.init_process_group("nccl", rank=rank, world_size=world, timeout=datetime.timedelta(seconds=300))
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    line, overflow = run_sharded(args, dev, rank, world)
    if rank == 0:
        os.write(real_stdout, (json.dumps(line) + "\n").encode())
    if world > 1:
        import torch.distributed as dist
        dist.destroy_process_group()
    if overflow:
        raise SystemExit("a sync-free render ran out of binning capacity: its image was incomplete, the timing is void")


if __name__ == "__main__":
    main()
