#!/usr/bin/env python
"""CF-3DGS global optimisation steps on a synthetic Tanks-scale sequence (BASELINE.json configs[2] and [3]).

Shape of the reference's GLOBAL STEPS (trainer/cf3dgs_trainer.py:324-372 with the per-iteration bookkeeping of
:144-170): a global model (~1M Gaussians, the union of per-frame pixel-aligned clouds) is optimised with Adam, each
iteration against ONE randomly chosen earlier frame rendered through that frame's (fixed) estimated camera;
loss = 0.8*L1 + 0.2*(1-SSIM); max_radii2D / add_densification_stats / optimizer step after every backward; every
`--densify-every` iterations the Gaussian count changes (scene/gaussian_model_cf.py:678-714 clones, splits and prunes
there; the surgery itself is outside the rasterizer path, so a deterministic clone / prune stand-in driven by the same
statistics changes the count here -- what is measured is how the operator copes with P and R jumping: binning
capacity, CUDA-graph re-capture, overflow reports).

Single GPU:
  python examples/cf3dgs_global.py [--frames 120] [--points-per-frame 8700] [--iters 600] [--impl ours|reference]
                                   [--graph] [--densify-every 100] [--psnr-every 100]
Batched global views over N GPUs: every rank holds a replica of the model and renders a different
training view per step; the gradients are written straight into one flat buffer and summed over NCCL (only the active SH
bands travel; at SH degree 2 and 3 the SH rows are reduce-scattered, stepped on 1/N of the rows per rank and all-gathered):
  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 examples/cf3dgs_global.py --batched ...
Prints one JSON line (rank 0).
"""
from __future__ import annotations

import argparse
import json
import math
import os
import sys
import time
import warnings

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "cf-3dgs_b200")); sys.path.insert(0, ROOT); sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))

import numpy as np
import torch
import torch.nn.functional as F

import gsr_synth as S
import cf3dgs_pair as CP

LRS = dict(xyz=0.00016, f_dc=0.0025, f_rest=0.000125, opacity=0.05, scaling=0.005, rotation=0.001)


def frame_pose(i, n):
    """Smooth camera trajectory: small yaw + sideways drift per frame (world-to-camera), the same total sweep for any
    number of frames."""
    s = 8.0 / max(n, 8)
    a = math.radians(0.5 * i * s)
    R = np.array([[math.cos(a), 0, math.sin(a)], [0, 1, 0], [-math.sin(a), 0, math.cos(a)]])
    t = np.array([0.03 * i * s, 0.0, 0.01 * i * s])
    return R, t


def default_config(**kw):
    cfg = dict(frames=8, points_per_frame=130_000, iters=200, width=960, height=540, sh_degree=3, impl="ours", loss=None,
               model=None, graph=False, batched=False, densify_every=0, psnr_every=0, profile=False, warmup=10, multicast="off", dc_plane="on", sh_up_every=0)
    cfg.update(kw)
    if cfg["model"] is None:
        cfg["model"] = "fused" if cfg["impl"] == "ours" else "torch"
    if cfg["loss"] is None:
        cfg["loss"] = "fused" if cfg["impl"] == "ours" else "torch"
    return argparse.Namespace(**cfg)


def _densify_stand_in(t, accum, denom, round_index):
    """Deterministic stand-in for densify_and_prune: clone the ~2 % of Gaussians with the largest mean screen-space
    gradient, drop the ~1.5 % with the lowest opacity (the fractions drift with the round so the count moves both
    ways).  `t`: dict of raw parameter tensors (+ optional Adam moments) indexed by Gaussian.  Returns the new dict."""
    P = t["xyz"].shape[0]
    g = accum / denom.clamp(min=1.0)
    n_clone = int(P * (0.010 + 0.004 * ((round_index * 7) % 5)))
    n_prune = int(P * (0.006 + 0.006 * ((round_index * 3) % 4)))
    clone_idx = torch.topk(g, n_clone).indices.sort().values
    keep = torch.ones(P, dtype=torch.bool, device=g.device)
    keep[torch.topk(-t["opacity_raw"].reshape(-1), n_prune).indices] = False
    keep_idx = torch.nonzero(keep).squeeze(1)
    out = {}
    for k, v in t.items():
        new = v[clone_idx].clone()
        if k in ("m", "v_"):     # cloned Gaussians start with empty moments, as densification_postfix gives them
            new.zero_()
        out[k] = torch.cat([v[keep_idx], new]).contiguous()
    return out


def run_global(args, dev, rank=0, world=1):
    """Runs the loop described above; returns the result dict (rank 0's view of it)."""
    import torch.distributed as dist
    W, H = args.width, args.height
    backend = CP.make_backend(args.impl, dev)
    loss_fn = CP.make_loss(args.loss)
    window = CP._gauss_window(device=dev)
    # ---- ground-truth scene, per-frame cameras and target images (identical on every rank: seeded) ----
    gt = S.make_gpix(W, H, seed=3, target_P=250_000)
    gt_t = {k: gt[k].to(dev) for k in ["means3D", "shs", "scales", "rotations"]}
    gt_opa = torch.full((gt_t["means3D"].shape[0], 1), 0.95, device=dev)
    eye, zero6 = torch.eye(4, device=dev), torch.zeros(6, device=dev)
    cams, targets = [], []
    for i in range(args.frames):
        R, t = frame_pose(i, args.frames)
        cam = S.make_camera(W, H, R=R, t=t)
        rs0 = backend["settings"](cam, 0)
        with torch.no_grad():
            targets.append(backend["render"](rs0, zero6, gt_t["means3D"], eye, gt_opa, shs=gt_t["shs"], scales=gt_t["scales"],
                                             rotations=gt_t["rotations"])[0].clamp(0, 1))
        cams.append(cam)
    # ---- global model: union of per-frame pixel-aligned clouds, moved into the world frame of camera 0 ----
    xs, ss, cs = [], [], []
    for i in range(args.frames):
        c = S.make_gpix(W, H, seed=3, target_P=args.points_per_frame)
        R, t = frame_pose(i, args.frames)
        Rt = torch.tensor(R, dtype=torch.float32); tt = torch.tensor(t, dtype=torch.float32)
        xs.append((c["means3D"] - tt) @ Rt)           # camera i -> world: R^T (p - t)
        ss.append(c["scales"]); cs.append(c["shs"])
    raw = dict(xyz=torch.cat(xs).to(dev), features=torch.cat(cs).to(dev).contiguous(),
               log_scales=torch.cat(ss).to(dev).log())
    P0 = raw["xyz"].shape[0]
    raw["rot_raw"] = torch.zeros(P0, 4, device=dev); raw["rot_raw"][:, 0] = 1
    raw["opacity_raw"] = torch.full((P0, 1), math.log(0.1 / 0.9), device=dev)
    settings = [backend["settings"](c, args.sh_degree) for c in cams]
    total_steps = args.iters + args.warmup
    order = np.random.RandomState(0).randint(0, args.frames, size=total_steps * max(world, 1))
    fused = args.model == "fused"
    if args.batched and not (fused and args.impl == "ours"):
        raise SystemExit("--batched needs --impl ours --model fused")
    if args.graph and not (fused and args.impl == "ours"):
        raise SystemExit("--graph needs --impl ours --model fused")
    if args.graph and args.sh_up_every:
        raise SystemExit("--sh-up-every changes the SH degree, which a captured graph has fixed")

    import gsr_parallel as GPar
    state = {"P": P0, "recaptures": 0, "recapture_s": 0.0, "overflow_warnings": 0, "densify_rounds": 0, "P_history": [P0]}

    # ---- the model in one of three forms ----
    def build_model(r, moments=None):
        if fused:
            import gsr_model
            m = gsr_model.FusedGaussianModel(r["xyz"], r["features"], r["opacity_raw"], r["log_scales"], r["rot_raw"], LRS,
                                             device_step=True)
            return m
        p = {k: v.clone().requires_grad_(True) for k, v in r.items() if k != "features"}
        p["f_dc"] = r["features"][:, :1].clone().requires_grad_(True)
        p["f_rest"] = r["features"][:, 1:].clone().requires_grad_(True)
        opt = torch.optim.Adam([{"params": [p["xyz"]], "lr": LRS["xyz"]}, {"params": [p["f_dc"]], "lr": LRS["f_dc"]},
                                {"params": [p["f_rest"]], "lr": LRS["f_rest"]}, {"params": [p["opacity_raw"]], "lr": LRS["opacity"]},
                                {"params": [p["log_scales"]], "lr": LRS["scaling"]}, {"params": [p["rot_raw"]], "lr": LRS["rotation"]}],
                               eps=1e-15, fused=True)
        stats = dict(accum=torch.zeros(r["xyz"].shape[0], device=dev), denom=torch.zeros(r["xyz"].shape[0], device=dev),
                     max_radii=torch.zeros(r["xyz"].shape[0], device=dev))
        return dict(p=p, opt=opt, stats=stats)

    model = build_model(raw)

    def raw_of(model):
        if fused:
            return dict(xyz=model.xyz, features=model.features, opacity_raw=model.opacity_raw, log_scales=model.log_scales,
                        rot_raw=model.rot_raw)
        p = model["p"]
        return dict(xyz=p["xyz"].detach(), features=torch.cat([p["f_dc"], p["f_rest"]], 1).detach(),
                    opacity_raw=p["opacity_raw"].detach(), log_scales=p["log_scales"].detach(), rot_raw=p["rot_raw"].detach())

    # ---- one iteration on view f ----
    def fused_iteration_frame(rs, target):
        m = model
        act = m.activated()
        x = m.xyz.detach().requires_grad_(True)
        sh = (m.shs(int(rs.sh_degree)) if args.dc_plane == "on" else m.features).detach().requires_grad_(True)     # degree 0: the dense coefficient-0 plane
        o, sc, r = act["opacities"].requires_grad_(True), act["scales"].requires_grad_(True), act["rotations"].requires_grad_(True)
        m2d = torch.zeros_like(x, requires_grad=True)
        import diff_gaussian_rasterization as D
        img, radii, _, _ = D.GaussianRasterizer(rs)(x, m2d, o, shs=sh, scales=sc, rotations=r)
        loss = loss_fn(img, target, window)
        loss.backward()
        m.step(x.grad, sh.grad, o.grad, sc.grad, r.grad, d_means2D=m2d.grad, radii=radii, sh_degree=int(rs.sh_degree))
        return loss

    def torch_iteration_frame(rs, target):
        p, opt, st = model["p"], model["opt"], model["stats"]
        # the reference keeps `viewspace_points` (means2D) to read its gradient (gaussian_renderer/__init__.py:55-60)
        out = backend["render"](rs, zero6, p["xyz"], eye, torch.sigmoid(p["opacity_raw"]), shs=torch.cat([p["f_dc"], p["f_rest"]], 1),
                                scales=torch.exp(p["log_scales"]), rotations=F.normalize(p["rot_raw"]))
        loss = loss_fn(out[0], target, window)
        loss.backward()
        with torch.no_grad():     # trainer/cf3dgs_trainer.py:147-152 (the reference arm has no means2D gradient handle here:
            vis = out[1] > 0 if out[1] is not None else None      # the statistics use the mean gradient of xyz instead)
            gnorm = p["xyz"].grad.norm(dim=1)
            if vis is None:
                vis = gnorm > 0
            st["accum"][vis] += gnorm[vis]; st["denom"][vis] += 1
        opt.step(); opt.zero_grad(set_to_none=True)
        return loss

    mcast = {"on": True, "auto": None, "off": False}[args.multicast]
    batched = GPar.BatchedViewsStep(model, multicast=mcast) if args.batched else None

    def view_of(k):
        return int(order[k * world + rank]) if args.batched else int(order[k])

    def settings_at(f, k):
        """The view's settings at iteration k: with --sh-up-every N the active SH degree starts at 0 and rises by one every
        N iterations up to --sh-degree (the reference's oneupSHdegree every 1000 iterations, trainer/cf3dgs_trainer.py:
        335-337 with scene/gaussian_model_cf.py:174-176)."""
        if not args.sh_up_every:
            return settings[f]
        return settings[f]._replace(sh_degree=min(args.sh_degree, max(k - args.warmup, 0) // args.sh_up_every))

    def eager_iteration(k):
        f = view_of(k)
        rs = settings_at(f, k)
        if batched is not None:
            return batched(rs, targets[f], loss_fn, window)
        return fused_iteration_frame(rs, targets[f]) if fused else torch_iteration_frame(rs, targets[f])

    # ---- CUDA-graph form: ONE graph for all views; the view's camera and target are gathered on the device ----
    graph = {"g": None}
    if args.graph:
        import diff_gaussian_rasterization as D
        views_all = torch.stack([s.viewmatrix for s in settings]); projs_all = torch.stack([s.projmatrix for s in settings])
        campos_all = torch.stack([s.campos for s in settings]); targets_all = torch.stack(targets)
        idx_dev = torch.zeros(1, dtype=torch.int64, device=dev)
        # the schedule of views lives on the device and the graph advances its own position in it: a host-side "next view"
        # word would be overwritten by the host, which runs ahead of the GPU, before the replay that should read it
        order_dev = torch.as_tensor(order[:total_steps], dtype=torch.int64, device=dev)
        pos_dev = torch.zeros(1, dtype=torch.int64, device=dev)
        cur = dict(view=torch.empty_like(settings[0].viewmatrix), proj=torch.empty_like(settings[0].projmatrix),
                   campos=torch.empty_like(settings[0].campos), target=torch.empty_like(targets[0]))
        rs_cur = settings[0]._replace(viewmatrix=cur["view"], projmatrix=cur["proj"], campos=cur["campos"])

        def graph_body(scheduled=False):
            if scheduled:
                idx_dev.copy_(order_dev.index_select(0, pos_dev))
                pos_dev.add_(1)
            cur["view"].copy_(views_all.index_select(0, idx_dev)[0]); cur["proj"].copy_(projs_all.index_select(0, idx_dev)[0])
            cur["campos"].copy_(campos_all.index_select(0, idx_dev)[0]); cur["target"].copy_(targets_all.index_select(0, idx_dev)[0])
            return fused_iteration_frame(rs_cur, cur["target"])

        def capture():
            """(Re-)capture after the Gaussian count has changed: every workspace size in the graph depends on P."""
            t0 = time.time()
            D.set_sync_free(True, slack=2.0)
            D._SyncFree.capacity.pop(dev.index, None)      # learn the capacity again: R moved with P
            side = torch.cuda.Stream()
            side.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(side), warnings.catch_warnings(record=True) as wlist:
                warnings.simplefilter("always")
                for f in range(0, args.frames, max(1, args.frames // 8)):    # a spread of views sets the capacity
                    idx_dev.fill_(f)
                    graph_body()
                D.sync_free_status()
                state["overflow_warnings"] += sum("exceeded the binning capacity" in str(w.message) for w in wlist)
            torch.cuda.current_stream().wait_stream(side)
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g):
                loss = graph_body(scheduled=True)
            torch.cuda.synchronize(dev)
            graph["g"], graph["loss"] = g, loss
            state["recaptures"] += 1
            state["recapture_s"] += time.time() - t0

        def graph_iteration(k):
            graph["g"].replay()            # renders view order[pos] and advances pos, all on the device
            return graph["loss"]

    def iteration(k):
        return graph_iteration(k) if args.graph else eager_iteration(k)

    # ---- the change of Gaussian count every `densify_every` iterations ----
    def densify(round_index):
        nonlocal model, batched
        r = raw_of(model)
        if fused:
            accum, denom, maxr = model.xyz_gradient_accum, model.denom, model.max_radii2D
            if args.batched:
                GPar.allreduce_densify_stats(accum, denom, maxr)      # sum, sum, max over the ranks' views
        else:
            accum, denom = model["stats"]["accum"], model["stats"]["denom"]
        new = _densify_stand_in({k: v.detach() for k, v in r.items()}, accum, denom, round_index)
        model = build_model(new)
        if args.batched:
            batched = GPar.BatchedViewsStep(model, multicast=mcast)
        state["P"] = new["xyz"].shape[0]
        state["P_history"].append(state["P"])
        state["densify_rounds"] += 1
        if args.graph:
            capture()

    def psnr_now():
        """Mean PSNR over 4 fixed views (render without gradient through the same backend)."""
        r = raw_of(model)
        vals = []
        with torch.no_grad():
            for f in range(0, args.frames, max(1, args.frames // 4))[:4]:
                img = backend["render"](settings[f], zero6, r["xyz"], eye, torch.sigmoid(r["opacity_raw"]), shs=r["features"],
                                        scales=torch.exp(r["log_scales"]), rotations=F.normalize(r["rot_raw"]))[0].clamp(0, 1)
                mse = ((img - targets[f]) ** 2).mean()
                vals.append(float(-10.0 * torch.log10(mse.clamp(min=1e-12))))
        return sum(vals) / len(vals)

    if args.graph:
        capture()
        state["recaptures"], state["recapture_s"] = 0, 0.0       # the first capture is set-up, not re-capture
    first = None
    for k in range(args.warmup):
        first = iteration(k)
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize(dev)
    psnr = [(0, psnr_now())] if args.psnr_every else []
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.time()
    e0.record()
    excluded = 0.0
    for it in range(args.iters):
        k = args.warmup + it
        loss = iteration(k)
        if args.densify_every and (it + 1) % args.densify_every == 0 and it + 1 < args.iters:
            densify(state["densify_rounds"])
        if args.psnr_every and (it + 1) % args.psnr_every == 0:
            torch.cuda.synchronize(dev)
            tp = time.time()
            psnr.append((it + 1, psnr_now()))
            torch.cuda.synchronize(dev)
            excluded += time.time() - tp
    e1.record()
    torch.cuda.synchronize(dev)
    dt = time.time() - t0 - excluded
    overflow = None
    if args.graph:
        import diff_gaussian_rasterization as D
        with warnings.catch_warnings(record=True) as wlist:
            warnings.simplefilter("always")
            st = D.sync_free_status()
        overflow = bool(st[1]) if st else False
        state["overflow_warnings"] += int(overflow)
        D.set_sync_free(False)
    identical = None
    if args.batched:
        r = raw_of(model)
        identical = GPar.assert_replicas_identical([r["xyz"], r["features"], r["opacity_raw"], r["log_scales"], r["rot_raw"]])
        if world > 1:
            tmax = torch.tensor([dt], device=dev)
            dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
            dt = float(tmax.item())
    deg = args.sh_degree
    line = {"workload": "cf3dgs_global_steps", "impl": args.impl, "loss": args.loss, "model": args.model, "P": P0,
            "P_final": state["P"], "frames": args.frames, "W": W, "H": H, "sh_degree": deg, "iters": args.iters,
            "iters_per_s": args.iters / dt, "ms_per_iter": dt / args.iters * 1e3, "first_loss": float(first.detach()) if first is not None else None,
            "last_loss": float(loss.detach()), "graph": bool(args.graph), "n_gpus": world, "batched_views": bool(args.batched)}
    if args.batched:
        line["views_per_s"] = world * args.iters / dt
        line["allreduce_floats_per_gaussian"] = 11 + 3 * (deg + 1) * (deg + 1)
        line["allreduce_bytes"] = 4 * (11 + 3 * (deg + 1) * (deg + 1)) * state["P"]
        line["sh_rows_sharded"] = bool(world > 1 and (deg + 1) * (deg + 1) >= GPar.BatchedViewsStep.SHARD_FROM_M)
        line["replicas_bit_identical"] = identical
        line["shared_step_multicast"] = batched.multicast_reason      # "in use", or why the NCCL form ran
    if args.densify_every:
        line["densify_every"] = args.densify_every
        line["densify_rounds"] = state["densify_rounds"]
        line["P_history"] = state["P_history"]
    if args.graph:
        line["graph_recaptures"] = state["recaptures"]
        line["graph_recapture_s"] = state["recapture_s"]
        line["binning_overflow"] = overflow
        line["overflow_warnings"] = state["overflow_warnings"]
    if psnr:
        line["psnr_curve"] = [[i, round(v, 3)] for i, v in psnr]
    if args.batched and identical is False:
        line["error"] = "replicas diverged"
    return line


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
    ap.add_argument("--graph", action="store_true",
                    help="sync-free forward + ONE CUDA graph holding the whole iteration for any view (ours, fused model)")
    ap.add_argument("--model", default=None, choices=["fused", "torch"],
                    help="fused activation/Adam/statistics kernels (default for ours) or torch.optim.Adam + autograd")
    ap.add_argument("--batched", action="store_true", help="batched global views: one view per rank per step, NCCL all-reduce")
    ap.add_argument("--multicast", default="off", choices=["on", "auto", "off"],
                    help="--batched: the optimiser step shared between the replicas over NVSwitch multicast (multimem loads / "
                         "stores) instead of NCCL collectives; auto = when the box supports it")
    ap.add_argument("--sh-up-every", type=int, default=0,
                    help="start at SH degree 0 and raise the active degree by one every N iterations up to --sh-degree (0 = fixed)")
    ap.add_argument("--dc-plane", default="on", choices=["on", "off"],
                    help="off: render SH degree 0 from the full (P,16,3) tensor instead of the model's dense coefficient-0 plane (A/B)")
    ap.add_argument("--densify-every", type=int, default=0, help="change the Gaussian count every N iterations (0 = never)")
    ap.add_argument("--psnr-every", type=int, default=0, help="record the PSNR over 4 fixed views every N iterations")
    ap.add_argument("--c3", action="store_true", help="BASELINE configs[2] at its stated shape: 120 frames, ~1M Gaussians, "
                                                      "count changing every 100 iterations, PSNR curve")
    a = ap.parse_args()
    if a.c3:
        a.frames, a.points_per_frame = 120, 8700
        a.densify_every = a.densify_every or 100
        a.psnr_every = a.psnr_every or 100
        a.iters = max(a.iters, 600)
    args = default_config(**{k: v for k, v in vars(a).items() if k != "c3"})
    real_stdout = os.dup(1); os.dup2(2, 1)
    world = int(os.environ.get("WORLD_SIZE", "1")); rank = int(os.environ.get("RANK", "0")); local = int(os.environ.get("LOCAL_RANK", "0"))
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    if world > 1:
        import datetime
        import torch.distributed as dist
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1"); os.environ.setdefault("MASTER_PORT", "29500")
        dist.init_process_group("nccl", rank=rank, world_size=world, timeout=datetime.timedelta(seconds=300))
        if not args.batched:
            raise SystemExit("more than one rank: pass --batched (independent replicas are what bench.py --gpus N measures)")
    line = run_global(args, dev, rank, world)
    if rank == 0:
        os.write(real_stdout, (json.dumps(line) + "\n").encode())
    if world > 1:
        import torch.distributed as dist
        dist.destroy_process_group()
    if line.get("error"):
        raise SystemExit(line["error"])


if __name__ == "__main__":
    main()
