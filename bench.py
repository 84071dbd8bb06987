#!/usr/bin/env python
"""bench.py -- headline benchmark of the B200 Gaussian-splatting rasterizer hot path.

Metric (BASELINE.json): forward+backward rendered frames/s at 1920x1080.
Workload at N=1: `gvol_1M_1080p_sh3` -- synthetic G-vol scene, 1,000,000 Gaussians, SH degree 3,
one fwd+bwd rasterization per step through the reference-shaped API (GaussianRasterizer autograd Function ->
C ABI -> sm_100a kernels), colour cotangent only (CF-3DGS default: depth/alpha losses off).
N>1 (one process per GPU, weak scaling): `value` is the sharded form north_star asks near-linear scaling of -- independent
frame pairs, every rank renders its own view of its own model replica with NO data-path collective;
value = N * frames / max-over-ranks time.  The same run measures the second half of BASELINE.json's metric and the form
that does communicate, as top-level numeric fields of the same JSON line:
  pose_opt_iters_per_s          CF-3DGS frame pairs sharded over the ranks (2 pairs per GPU, examples/cf3dgs_pair.py),
  gaussian_fit_iters_per_s      the local Gaussian fits of the same pairs,
  global_views_iters_per_s_sh0 / _sh3 and global_views_per_s_sh0 / _sh3
                                batched global views (8e (ii), examples/cf3dgs_global.py --batched): replicated 1.04 M-
                                Gaussian model, one view per rank per optimiser step, gradients written into one flat
                                buffer and all-reduced over NCCL (only the active SH bands), fused Adam on every rank.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

`--impl reference` times the UNMODIFIED reference rasterizer: its own CUDA build from oracle/_ref (compiled by
oracle/build_ref.py from the sources under /root/reference) when that exists and a GPU is present, otherwise the
CPU oracle port on a bounded sample.  The oracle is only ever the checker / baseline here, never the product path.
"""
from __future__ import annotations

import argparse
import json
import math
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(ROOT, "cf-3dgs_b200"))
sys.path.insert(0, ROOT)

import numpy as np
import torch

WORKLOAD = dict(name="gvol_1M_1080p_sh3", P=1_000_000, W=1920, H=1080, seed=5, sh_degree=3)
PARAM_KEYS = ["means3D", "opacities", "shs", "scales", "rotations"]
# DRAM bytes per launch of the kernels that can dominate, from the committed ncu capture of this workload
NCU_TRAFFIC_BYTES = {"render_backward": 150.4e6 + 17.9e6, "render_forward": 66.8e6 + 29.1e6,
                     "preprocess": 253.4e6 + 40.1e6, "preprocess_backward": 284.8e6 + 201.8e6}
# warp instructions per launch of the two blend kernels on this workload (smsp__inst_executed.sum, same ncu capture):
# what their issue-slot utilisation is computed from
NCU_WARP_INSTRUCTIONS = {"render_backward": 659.5e6, "render_forward": 391.8e6}
NCU_SOURCE = "profiles/r02_ncu_step_1M_summary.txt"


def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            d = json.load(f)
        return float(d.get("hbm_gbs", 6650.0)), "measured"
    return 6650.0, "fallback"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md)."""
    Q = "clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
        "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index):
        self.index = index
        self.rows = []       # (host time the row was read, row)
        self.proc = None
        self.marks = {}

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "20"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append((time.time(), line.strip()))

    def mark(self, name):
        self.marks[name] = time.time()

    def stop(self):
        """Clocks over [load_begin, load_end] -- the timed K steps plus the same step repeated right after them for long
        enough to collect samples -- with the number of samples that fell inside the timed K steps themselves."""
        if self.proc is None:
            return None
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        lo, hi = self.marks.get("load_begin", 0.0), self.marks.get("load_end", float("inf"))
        t0, t1 = self.marks.get("timed_begin", lo), self.marks.get("timed_end", hi)
        sm, mx, reasons, inside = [], [], set(), 0
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ts, r in self.rows:
            if not (lo <= ts <= hi + 0.03):
                continue
            parts = [x.strip() for x in r.split(",")]
            if len(parts) < 7:
                continue
            try:
                sm.append(float(parts[0])); mx.append(float(parts[1]))
            except ValueError:
                continue
            inside += int(t0 <= ts <= t1 + 0.03)
            for n, v in zip(names, parts[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        if not sm:
            return None
        return {"sm_mhz": statistics.median(sm), "sm_max_mhz": max(mx), "reasons": sorted(reasons), "samples": len(sm),
                "samples_inside_timed_steps": inside,
                "window": "the timed steps plus the same step repeated for >= 0.5 s directly after them (under load throughout)"}


def dist_setup(n_gpus, init=True):
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not init:      # the reference arm runs on rank 0 alone and never communicates
        return 1, rank, local
    if world > 1:
        import torch.distributed as dist
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29500")
        backend = "nccl" if torch.cuda.is_available() else "gloo"
        import datetime
        dist.init_process_group(backend=backend, rank=rank, world_size=world, timeout=datetime.timedelta(seconds=120))
    return world, rank, local


def rank_camera(S, W, H, rank, world):
    """View of rank r: yaw offset around the canonical camera so every rank renders different pixels."""
    if world == 1:
        return S.make_camera(W, H)
    yaw = (rank - (world - 1) / 2.0) * 1.5
    a = math.radians(yaw)
    R = np.array([[math.cos(a), 0, math.sin(a)], [0, 1, 0], [-math.sin(a), 0, math.cos(a)]])
    return S.make_camera(W, H, R=R, t=np.array([0.02 * (rank - (world - 1) / 2.0), 0.0, 0.0]))


def timed(step, K, warm, world, dev, clock_index=None):
    import torch.distributed as dist
    for _ in range(warm):
        step()
    sampler = ClockSampler(clock_index) if clock_index is not None else None
    if sampler:
        sampler.start()
        time.sleep(0.05)
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize(dev)
    if sampler:
        sampler.mark("load_begin"); sampler.mark("timed_begin")
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(K):
        step()
    e1.record()
    torch.cuda.synchronize(dev)
    if sampler:
        sampler.mark("timed_end")
    if world > 1:
        dist.barrier()
    ms = e0.elapsed_time(e1)
    clocks = None
    if sampler:
        # nvidia-smi samples every 20 ms and the K timed steps may last less than that: keep the same load on the GPU
        # until the sampler has seen it for half a second, so the clock record speaks about this workload
        t_end = time.time() + max(0.5 - ms / 1000.0, 0.1)
        while time.time() < t_end:
            for _ in range(10):
                step()
            torch.cuda.synchronize(dev)
        sampler.mark("load_end")
        clocks = sampler.stop()
    if world > 1:
        t = torch.tensor([ms], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    return ms, clocks


def cpu_oracle_baseline(S, wl, cam, sample_frac=1.0):
    """C oracle port, OpenMP over all host cores, one fwd+bwd frame of the same workload."""
    from oracle import oracle as O
    sc = S.make_gvol(wl["P"], wl["W"], wl["H"], seed=wl["seed"], sh_degree=wl["sh_degree"])
    ct = S.cotangents(wl["W"], wl["H"], seed=1, dense=False)
    cores = os.cpu_count() or 1
    t0 = time.time()
    st = O.forward(sc["means3D"].numpy(), sc["opacities"].numpy(), shs=sc["shs"].numpy(), scales=sc["scales"].numpy(),
                   rotations=sc["rotations"].numpy(), viewmatrix=cam["viewmatrix"].numpy(),
                   projmatrix=cam["projmatrix"].numpy(), campos=cam["campos"].numpy(), bg=np.zeros(3, np.float32),
                   W=wl["W"], H=wl["H"], tanfovx=cam["tanfovx"], tanfovy=cam["tanfovy"], sh_degree=wl["sh_degree"])
    O.backward(st, ct["grad_color"].numpy())
    dt = time.time() - t0
    return {"value": 1.0 / dt, "unit": "frames/s", "cores": cores, "kind": "port",
            "sample": f"1 full fwd+bwd frame of {wl['name']} (C oracle, OpenMP {cores} threads, {dt:.2f} s)"}


def torch_cpu_baseline(S):
    """north_star's "PyTorch-CPU evaluation of the same rasterization equations": oracle/torch_oracle.py (plain tensor ops,
    autograd backward) on the box's host cores, on a bounded sample -- BASELINE.json configs[0], 100k Gaussians at 640x360."""
    from oracle import torch_oracle as TO
    P, W, H = 100_000, 640, 360
    cam = S.make_camera(W, H)
    sc = S.make_gvol(P, W, H, seed=1, sh_degree=3)
    lv = {k: sc[k].clone().requires_grad_(True) for k in PARAM_KEYS}
    t0 = time.time()
    out = TO.forward(lv["means3D"], lv["opacities"], shs=lv["shs"], scales=lv["scales"], rotations=lv["rotations"],
                     viewmatrix=cam["viewmatrix"], projmatrix=cam["projmatrix"], campos=cam["campos"], bg=torch.zeros(3),
                     W=W, H=H, tanfovx=cam["tanfovx"], tanfovy=cam["tanfovy"], sh_degree=3)
    out["color"].sum().backward()
    dt = time.time() - t0
    return {"value": 1.0 / dt, "unit": "frames/s", "cores": torch.get_num_threads(), "kind": "port",
            "sample": f"1 fwd+bwd frame of 100k Gaussians at 640x360 (PyTorch CPU tensor ops + autograd, {dt:.1f} s); "
                      "not the 1M/1080p workload"}


def run_loops(dev, rank, world):
    """The second half of BASELINE.json's metric ("pose-opt iters/s at 1/2/4/8 GPU") and the batched-global-views form,
    measured in the same process right after the frames/s numbers (all ranks take part).  Returns a flat dict of numeric
    top-level fields for the JSON line."""
    sys.path.insert(0, os.path.join(ROOT, "examples"))
    import cf3dgs_pair as CP
    import cf3dgs_global as CG
    import diff_gaussian_rasterization as D
    out = {}
    D.set_sync_free(False)
    D._SyncFree.capacity.clear()      # the 1080p capacity of the frames/s part would oversize every launch of the 540p loops
    # (i) CF-3DGS frame pairs, 2 per GPU: local Gaussian fit (200 iterations) + pose fit (300 iterations, the reference's
    # count, trainer/cf3dgs_trainer.py:262-285), each loop one replayed CUDA graph; no data-path collective
    pargs = CP.default_args(pairs=2 * world, fit_iters=200, pose_iters=300, graph=True)
    line, overflow = CP.run_sharded(pargs, dev, rank, world)
    D.sync_free_status(); D.set_sync_free(False)
    D._SyncFree.capacity.clear()
    if rank == 0:
        out["pose_opt_iters_per_s"] = line["pose_opt_iters_per_s_aggregate"]
        out["gaussian_fit_iters_per_s"] = line["gaussian_fit_iters_per_s_aggregate"]
        out["pair_loop"] = {"pairs": line["pairs"], "P": line["P"], "W": line["W"], "H": line["H"], "fit_iters": 200,
                            "pose_iters": 300, "max_pose_err_t": line["max_pose_err_t"],
                            "max_pose_err_r_deg": line["max_pose_err_r_deg"], "binning_overflow": bool(overflow),
                            "what": "examples/cf3dgs_pair.py --graph: 518 k-Gaussian local models at 960x540, 2 pairs per GPU, "
                                    "pairs sharded over the ranks, one gather of the relative poses at the end"}
    # (ii) batched global views: replicated 1.04 M-Gaussian model, one view per rank per optimiser step, at SH degree 0 and
    # 3.  With several ranks the optimiser step is SHARED over NVSwitch multicast when the box can (every rank sums the
    # replicas' gradients of its 1/N of the rows with multimem.ld_reduce, steps them, and writes the new parameters to every
    # replica with multimem.st); the NCCL form (all-reduce / reduce-scatter + all-gather) is timed beside it.
    for deg in (0, 3):
        gargs = CG.default_config(frames=8, points_per_frame=130_000, iters=100, sh_degree=deg, batched=True, warmup=10,
                                  multicast="auto" if world > 1 else "off")
        g = CG.run_global(gargs, dev, rank, world)
        shared = g.get("shared_step_multicast") == "in use"      # the same answer on every rank (gsr_parallel agrees on it)
        if rank == 0:
            out[f"global_views_iters_per_s_sh{deg}"] = g["iters_per_s"]
            out[f"global_views_per_s_sh{deg}"] = g["views_per_s"]
            out.setdefault("global_views_loop", {})[f"sh{deg}"] = {
                "P": g["P"], "gradient_bytes_per_replica": g["allreduce_bytes"] if world > 1 else 0,
                "floats_per_gaussian": g["allreduce_floats_per_gaussian"], "replicas_bit_identical": g["replicas_bit_identical"],
                "ms_per_iter": g["ms_per_iter"], "last_loss": g["last_loss"],
                "step": ("shared over NVSwitch multicast" if shared else
                         "NCCL" + (f" ({g.get('shared_step_multicast')})" if world > 1 else " not needed: one rank"))}
        done_nccl = not shared
        if world > 1 and not done_nccl:
            gargs = CG.default_config(frames=8, points_per_frame=130_000, iters=100, sh_degree=deg, batched=True, warmup=10,
                                      multicast="off")
            g = CG.run_global(gargs, dev, rank, world)
            if rank == 0:
                out[f"global_views_nccl_iters_per_s_sh{deg}"] = g["iters_per_s"]
                out["global_views_loop"][f"sh{deg}"]["nccl_form_ms_per_iter"] = g["ms_per_iter"]
                out["global_views_loop"][f"sh{deg}"]["nccl_form_replicas_bit_identical"] = g["replicas_bit_identical"]
    if rank == 0:
        out["global_views_loop"]["what"] = ("examples/cf3dgs_global.py --batched: full iteration (activations, render, fused "
                                            "L1+SSIM, backward into the flat buffer, the optimiser step over the sum of the "
                                            "replicas' gradients, densification statistics), 960x540, 8 training views; the "
                                            "NCCL form is one all-reduce at SH degree 0 and, at degree 3, all-reduce of the "
                                            "geometry block + reduce-scatter / Adam on 1/N of the SH rows / all-gather")
    return out


_REAL_STDOUT = None


def emit(line: dict) -> None:
    """The ONE JSON line goes to the process's original stdout; everything else (NCCL banners, warnings) was
    re-routed to stderr at start-up so it cannot pollute it."""
    data = (json.dumps(line) + "\n").encode()
    if _REAL_STDOUT is None:
        sys.stdout.write(data.decode()); sys.stdout.flush()
    else:
        os.write(_REAL_STDOUT, data)


def main():
    global _REAL_STDOUT
    sys.stdout.flush()
    _REAL_STDOUT = os.dup(1)
    os.dup2(2, 1)
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-graph", action="store_true", help="run the end-to-end step eagerly instead of replaying a CUDA graph")
    ap.add_argument("--no-loops", action="store_true", help="skip the CF-3DGS pair / global-views loops (pose_opt_iters_per_s ...)")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3)
    world, rank, local = dist_setup(args.gpus, init=(args.impl == "ours"))
    import gsr_synth as S
    wl = dict(WORKLOAD)
    W, H, P = wl["W"], wl["H"], wl["P"]

    have_gpu = torch.cuda.is_available()
    if args.impl == "reference":
        if rank != 0:
            return 0
        from oracle import ref_cuda as RC
        if not (have_gpu and RC.available()):
            cam = S.make_camera(W, H)
            cb = cpu_oracle_baseline(S, wl, cam)
            line = {"impl": "reference", "metric": "fwd+bwd rendered frames/s at 1080p", "value": cb["value"],
                    "unit": "frames/s", "n_gpus": args.gpus, "steps": 1, "warmup": 0, "ms_per_step": 1000.0 / cb["value"],
                    "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
                    "config": {"workload": wl["name"], "P": P, "W": W, "H": H, "sh_degree": wl["sh_degree"]},
                    "cpu_baseline": cb, "gpu_launches": 0,
                    "e2e": {"value": cb["value"], "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
            emit(line)
            return 0

    if not have_gpu:
        raise SystemExit("bench.py needs a CUDA device: the product path has no CPU fallback")
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    cam = rank_camera(S, W, H, rank, world)
    sc = S.make_gvol(P, W, H, seed=wl["seed"], sh_degree=wl["sh_degree"])
    params = {k: sc[k].to(dev) for k in PARAM_KEYS}
    bg = torch.zeros(3, device=dev)
    view, proj, campos = cam["viewmatrix"].to(dev), cam["projmatrix"].to(dev), cam["campos"].to(dev)
    gc = S.cotangents(W, H, seed=1, dense=False)["grad_color"].to(dev)
    flat_grad = torch.empty(P * 59, device=dev) if world > 1 else None
    state = {}

    peaks = None
    if args.impl == "ours":
        import diff_gaussian_rasterization as D
        import gsr_loss
        lib = D.lib
        if rank == 0:
            # measured ceilings for the blend kernels (FFMA, FFMA2, MUFU.EX2, LDS), before anything is timed
            import ctypes
            buf4 = (ctypes.c_double * 4)()
            if lib.gsr_measure_peaks(buf4, torch.cuda.current_stream().cuda_stream) == 0:
                peaks = {"ffma_tflops": buf4[0] / 1e12, "ffma2_tflops": buf4[1] / 1e12, "mufu_ex2_gops": buf4[2] / 1e9,
                         "lds_gbs": buf4[3] / 1e9, "issue_gwarp_inst_per_s": buf4[0] / 64.0 / 1e9,
                         "how": "gsr_measure_peaks (csrc/gsr_peaks.cu): 8 CTAs x 256 threads per SM, best of 3 after a warm-up "
                                "run, CUDA events; issue ceiling = FFMA warp instructions per second"}
        rs = D.GaussianRasterizationSettings(image_height=H, image_width=W, tanfovx=cam["tanfovx"], tanfovy=cam["tanfovy"],
                                             bg=bg, scale_modifier=1.0, viewmatrix=view, projmatrix=proj,
                                             sh_degree=wl["sh_degree"], campos=campos, prefiltered=False, debug=False)

        def render(leaves, m2d, settings=rs):
            return D.GaussianRasterizer(settings)(leaves["means3D"], m2d, leaves["opacities"], shs=leaves["shs"],
                                                  scales=leaves["scales"], rotations=leaves["rotations"])
    else:
        from oracle import ref_cuda as RC

        class _RefFn(torch.autograd.Function):
            @staticmethod
            def forward(ctx, means3D, opac, shs, scales, rots, v, pj, cp):
                st = RC.forward(means3D, opac, shs=shs, scales=scales, rotations=rots, viewmatrix=v, projmatrix=pj,
                                campos=cp, bg=bg, W=W, H=H, tanfovx=cam["tanfovx"], tanfovy=cam["tanfovy"],
                                sh_degree=wl["sh_degree"])
                # as the reference's own wrapper: state through save_for_backward (outputs kept on ctx directly would form
                # a reference cycle and leave the workspaces to the garbage collector)
                ctx.save_for_backward(st["radii"], st["geom"], st["binning"], st["img"], st["alpha"])
                ctx.meta = (st["num_rendered"], st["_args"])
                return st["color"], st["depth"], st["alpha"]

            @staticmethod
            def backward(ctx, gcol, gdep, galp):
                radii, geom, binning, img, alpha = ctx.saved_tensors
                st = dict(radii=radii, geom=geom, binning=binning, img=img, alpha=alpha, num_rendered=ctx.meta[0], _args=ctx.meta[1])
                g = RC.backward(st, gcol.contiguous(), gdep.contiguous(), galp.contiguous())
                return g["means3D"], g["opacities"], g["sh"], g["scales"], g["rotations"], None, None, None

        def render(leaves, m2d, settings=None):
            v, pj, cp = (view, proj, campos) if settings is None else settings
            c, d, a = _RefFn.apply(leaves["means3D"], leaves["opacities"], leaves["shs"], leaves["scales"],
                                   leaves["rotations"], v, pj, cp)
            return c, None, d, a

    def step(comm=True):
        leaves = {k: v.detach().requires_grad_(True) for k, v in params.items()}
        m2d = torch.zeros_like(leaves["means3D"], requires_grad=True) if args.impl == "ours" else None
        out = render(leaves, m2d)
        out[0].backward(gc)
        if world > 1 and comm:
            import torch.distributed as dist
            off = 0
            for k in PARAM_KEYS:
                g = leaves[k].grad
                flat_grad[off:off + g.numel()].copy_(g.reshape(-1))
                off += g.numel()
            dist.all_reduce(flat_grad)
        state["leaves"] = leaves

    # ---- device-resident throughput ----
    ms, clocks = timed(lambda: step(comm=False), args.steps, args.warmup, world, dev,
                       clock_index=local if rank == 0 else None)
    value = world * args.steps / (ms / 1000.0)
    ar_ms = None
    if world > 1:   # batched views + gradient all-reduce, same step plus the collective
        ar_ms, _ = timed(lambda: step(comm=True), args.steps, 3, world, dev)

    # ---- batched views on two streams (SURVEY 8e (ii) on one GPU): the binning stages of one view are latency-bound and
    # leave SMs idle that the blend kernels of another view can use.  Reported beside `value`, not instead of it. ----
    two_stream = None
    if world == 1 and args.impl == "ours":
        streams = [torch.cuda.Stream(device=dev) for _ in range(2)]
        for st_ in streams:
            st_.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(st_):
                for _ in range(3):
                    step(comm=False)
        torch.cuda.synchronize(dev)
        t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0.record()
        for st_ in streams:
            st_.wait_event(t0)
        for _ in range(args.steps):
            for st_ in streams:
                with torch.cuda.stream(st_):
                    step(comm=False)
        for st_ in streams:
            torch.cuda.current_stream().wait_stream(st_)
        t1.record()
        torch.cuda.synchronize(dev)
        two_stream = 2 * args.steps / (t0.elapsed_time(t1) / 1000.0)

    # ---- end to end: per step the camera + target image come from pinned host memory, the loss goes back ----
    tgt_host = torch.rand(3, H, W).pin_memory()
    view_h, proj_h, cam_h = cam["viewmatrix"].pin_memory(), cam["projmatrix"].pin_memory(), cam["campos"].pin_memory()
    # double-buffered device copies: step i+1's inputs travel on a copy stream while step i computes
    tgt_dev = [torch.empty(3, H, W, device=dev) for _ in range(2)]
    view_d = [torch.empty_like(view) for _ in range(2)]
    proj_d = [torch.empty_like(proj) for _ in range(2)]
    cam_d = [torch.empty_like(campos) for _ in range(2)]
    loss_host = torch.zeros(1).pin_memory()
    h2d = tgt_host.numel() * 4 + (view_h.numel() + proj_h.numel() + cam_h.numel()) * 4
    d2h = 4
    copy_stream = torch.cuda.Stream(device=dev)
    ready = [torch.cuda.Event() for _ in range(2)]     # inputs of slot landed
    consumed = [torch.cuda.Event() for _ in range(2)]  # compute finished reading slot
    counter = {"i": 0, "primed": False}

    def upload(slot):
        with torch.cuda.stream(copy_stream):
            copy_stream.wait_event(consumed[slot])
            tgt_dev[slot].copy_(tgt_host, non_blocking=True)
            view_d[slot].copy_(view_h, non_blocking=True); proj_d[slot].copy_(proj_h, non_blocking=True)
            cam_d[slot].copy_(cam_h, non_blocking=True)
            ready[slot].record(copy_stream)

    # The compute of a step (render, L1 loss, backward) is captured once as a CUDA graph per input slot -- the sync-free
    # forward makes that possible -- and replayed; uploads, the loss read-back and the per-step synchronisation stay
    # outside the graph.  The reference arm runs eagerly (it synchronises inside its forward).
    use_graph = (args.impl == "ours" and not args.no_graph)
    graphs = [None, None]
    loss_dev = [torch.zeros((), device=dev) for _ in range(2)]
    static_leaves = None

    def compute(slot):
        leaves = static_leaves if static_leaves is not None else {k: v.detach().requires_grad_(True) for k, v in params.items()}
        if args.impl == "ours":
            st2 = rs._replace(viewmatrix=view_d[slot], projmatrix=proj_d[slot], campos=cam_d[slot])
            m2d = torch.zeros_like(leaves["means3D"], requires_grad=True)
        else:
            st2, m2d = (view_d[slot], proj_d[slot], cam_d[slot]), None
        color = render(leaves, m2d, st2)[0]
        if args.impl == "ours":       # the library's fused loss at lambda_dssim = 0: the plain L1 term, two small kernels
            loss = gsr_loss.photometric_loss(color, tgt_dev[slot], 0.0, clamp=False)
        else:
            loss = (color - tgt_dev[slot]).abs().mean()
        loss.backward()
        loss_dev[slot].copy_(loss.detach())
        return leaves

    if use_graph:
        D.set_sync_free(True, slack=1.5)
        static_leaves = {k: v.detach().clone().requires_grad_(True) for k, v in params.items()}
        for slot in range(2):                        # real cameras / targets in both slots before anything is rendered
            tgt_dev[slot].copy_(tgt_host); view_d[slot].copy_(view_h); proj_d[slot].copy_(proj_h); cam_d[slot].copy_(cam_h)
        compute(0)                                   # learns the binning capacity through the synchronising path
        side = torch.cuda.Stream(device=dev)
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            compute(0); compute(1)
        torch.cuda.current_stream().wait_stream(side)
        for slot in range(2):
            for v in static_leaves.values():
                v.grad = None
            graphs[slot] = torch.cuda.CUDAGraph()
            with torch.cuda.graph(graphs[slot]):
                compute(slot)

    def e2e_step():
        cur = torch.cuda.current_stream()
        slot = counter["i"] & 1
        if not counter["primed"]:
            consumed[0].record(cur); consumed[1].record(cur)
            upload(slot)
            counter["primed"] = True
        upload(slot ^ 1)                      # next step's inputs, overlapped with this step's kernels
        cur.wait_event(ready[slot])
        if use_graph:
            graphs[slot].replay()
        else:
            leaves = compute(slot)
        consumed[slot].record(cur)
        loss_host.copy_(loss_dev[slot].reshape(1), non_blocking=True)
        cur.synchronize()   # the caller reads the loss every step
        if os.environ.get("GSR_BENCH_DEBUG"):
            print("e2e step", counter["i"], "slot", slot, "loss", float(loss_host[0]), file=sys.stderr)
        counter["i"] += 1

    e2e_ms, _ = timed(e2e_step, args.steps, 3, world, dev)
    e2e_value = world * args.steps / (e2e_ms / 1000.0)
    if use_graph:
        ovf = D.sync_free_status()
        D.set_sync_free(False)
        if ovf is not None and ovf[1]:
            raise SystemExit("bench: sync-free binning capacity overflowed; the e2e number would be invalid")

    loops = {}
    if args.impl == "ours" and not args.no_loops:
        try:
            loops = run_loops(dev, rank, world)
        except Exception as ex:      # the loop numbers must never take the headline line down with them
            if world > 1:
                raise                # (with several ranks a one-sided failure would leave the others in a collective)
            loops = {"loops_error": f"{type(ex).__name__}: {ex}"}
    if rank != 0:
        return 0
    line = {"metric": "fwd+bwd rendered frames/s at 1080p", "value": value, "unit": "frames/s", "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms / args.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": wl["name"], "P": P, "W": W, "H": H, "sh_degree": wl["sh_degree"],
                       "cotangents": "colour only (CF-3DGS default)", "parallelism": f"dp{world}: independent frame pairs per rank, no data-path collective" if world > 1 else "dp1",
                       "l2": "per-step working set ~0.9 GB > 126 MB L2, no explicit flush"},
            "gaussians_per_s": value * P, "clocks": clocks,
            "e2e": {"value": e2e_value, "unit": "frames/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "what": "per step: camera + target image H2D from pinned memory (double-buffered on a copy stream), "
                            "render + L1 loss (" + ("gsr_loss, lambda_dssim = 0" if args.impl == "ours" else "PyTorch ops") + ") + backward (" + ("one CUDA graph replay, sync-free forward" if use_graph else "eager") +
                            "), loss D2H + sync"}}
    line.update(loops)
    if peaks is not None:
        line["measured_peaks"] = peaks
    if two_stream is not None:
        line["two_stream_variant"] = {"what": "two views in flight on two CUDA streams (batched views on one GPU): one view's "
                                              "latency-bound binning overlaps the other's blend kernels",
                                      "value": two_stream, "unit": "frames/s"}
    if ar_ms is not None:
        line["allreduce_variant"] = {"what": "same step + one NCCL all-reduce(sum) of the flat 59-float/Gaussian gradient "
                                             "(batched global views, SURVEY 8e (ii))",
                                     "value": world * args.steps / (ar_ms / 1000.0), "unit": "frames/s",
                                     "ms_per_step": ar_ms / args.steps, "allreduce_bytes": P * 59 * 4}
    if args.impl == "reference":
        line["impl"] = "reference"
        line["n_gpus"] = args.gpus          # the run it belongs to; the reference itself runs on rank 0 alone
        line["ranks_used"] = 1
        line["comparable_n"] = 1            # one GPU's worth of work: compare with the N = 1 line of the other arm only
        line["gpu_launches"] = 0
        line["cpu_baseline"] = {"value": value, "unit": "frames/s", "cores": 0, "kind": "reference",
                                "sample": "the reference's own CUDA rasterizer (oracle/_ref, unmodified sources, sm_100a) "
                                          "on this GPU: it has no CPU implementation"}
        emit(line)
        return 0

    # ---- per-stage timing, work counters, roofline of the dominant kernel ----
    hbm_peak, peak_kind = load_peaks()
    n_st = lib.gsr_profile_count()
    names = [lib.gsr_profile_name(i).decode() for i in range(n_st)]
    import ctypes
    acc = np.zeros(n_st)
    reps = 5
    lib.gsr_profile_enable(1)
    for _ in range(reps):
        step(comm=False)   # rank 0 only from here on: no collective
        buf = (ctypes.c_float * n_st)()
        lib.gsr_profile_read(buf, n_st)
        acc += np.maximum(np.array(buf[:]), 0.0)
    lib.gsr_profile_enable(0)
    stage_ms = {n: float(a / reps) for n, a in zip(names, acc)}
    # work counters from one more forward
    leaves = {k: v.detach().requires_grad_(True) for k, v in params.items()}
    out = render(leaves, torch.zeros_like(leaves["means3D"], requires_grad=True))
    fn = out[0].grad_fn
    geom, binning, img = fn.saved_tensors[7], fn.saved_tensors[8], fn.saved_tensors[9]
    R = fn.num_rendered
    counts = torch.zeros(4, dtype=torch.int64, device=dev)
    lib.gsr_debug_pair_counts(P, W, H, R, geom.data_ptr(), binning.data_ptr(), img.data_ptr(), counts.data_ptr(),
                              torch.cuda.current_stream().cuda_stream)
    in_range, blended, ref_pairs, tile_rej = [int(x) for x in counts.tolist()]
    P_vis = int((out[1] > 0).sum())
    tiles = ((W + 15) // 16) * ((H + 15) // 16)
    sm_mhz = (clocks or {}).get("sm_mhz") or 1500.0
    n_sm = torch.cuda.get_device_properties(dev).multi_processor_count
    if peaks is not None:
        fp32_peak = peaks["ffma_tflops"]
        fp32_source = "measured FFMA peak of this GPU in this run (gsr_measure_peaks, csrc/gsr_peaks.cu)"
    else:
        fp32_peak = n_sm * 128 * 2 * sm_mhz * 1e6 / 1e12
        fp32_source = f"derived FP32 FMA peak {n_sm} SMs x 128 lanes x 2 x {sm_mhz:.0f} MHz (the microbenchmark did not run)"
    # algorithmic work: bytes for the HBM-bound stages, flops for the blend kernels
    alg = {
        "preprocess": ("hbm", (P_vis * 311 + (P - P_vis) * 20) / 1e9),
        "depth_sort": ("hbm", P * 16 / 1e9),
        "offsets_scan": ("hbm", P * 8 / 1e9),
        "emit": ("hbm", (P * 20 + R * 12) / 1e9),
        "tile_sort": ("hbm", R * 24 / 1e9),      # (the tile ranges are written by its last pass)
        "render_forward": ("fp32", (in_range * 24 + blended * 16) / 1e12),
        "render_backward": ("fp32", (blended * 75 + (in_range - blended) * 24) / 1e12),
        "preprocess_backward": ("hbm", P_vis * 627 / 1e9),
    }
    stages = {}
    for n, (bound, work) in alg.items():
        t = stage_ms.get(n, 0.0)
        if t <= 0:
            continue
        ach = work / (t / 1000.0)
        peak = hbm_peak if bound == "hbm" else fp32_peak
        stages[n] = {"ms": t, "bound": bound, "achieved": ach, "unit": "GB/s" if bound == "hbm" else "TFLOP/s",
                     "frac": ach / peak}
        if n in NCU_WARP_INSTRUCTIONS and peaks is not None:
            # an issue-bound kernel is also judged against the issue-slot ceiling: warp instructions per second
            rate = NCU_WARP_INSTRUCTIONS[n] / (t / 1000.0) / 1e9
            stages[n]["issue"] = {"achieved_gwarp_inst_per_s": rate, "peak_gwarp_inst_per_s": peaks["issue_gwarp_inst_per_s"],
                                  "frac": rate / peaks["issue_gwarp_inst_per_s"],
                                  "warp_instructions_per_launch": NCU_WARP_INSTRUCTIONS[n], "source": NCU_SOURCE}
    dom = max(stages, key=lambda k: stages[k]["ms"])
    d = stages[dom]
    line["roofline"] = {"kernel": dom, "bound": d["bound"], "achieved": d["achieved"],
                        "peak": hbm_peak if d["bound"] == "hbm" else fp32_peak, "unit": d["unit"], "frac": d["frac"],
                        "traffic": NCU_TRAFFIC_BYTES.get(dom), "ms": d["ms"],
                        "traffic_source": "dram__bytes_read.sum + dram__bytes_write.sum per launch, ncu --set full, " +
                                          NCU_SOURCE + " (same workload)",
                        "peak_source": (f"{peak_kind} HBM copy" if d["bound"] == "hbm" else fp32_source)}
    if "issue" in d:
        line["roofline"]["issue_slot_frac"] = d["issue"]["frac"]
    line["stages"] = stages
    line["work"] = {"num_rendered": R, "visible": P_vis, "pairs_in_range": in_range, "pairs_blended": blended,
                    "reference_pairs_no_early_exit": ref_pairs, "tile_rejected_entries": tile_rej}
    depth_passes, tile_passes = 4, (max(1, (tiles - 1).bit_length()) + 7) // 8
    # preprocess, 3 kernels per radix pass, partial sums, emit, forward blend, backward blend, per-Gaussian backward
    line["gpu_launches"] = args.steps * (1 + 3 * depth_passes + 1 + 1 + 3 * tile_passes + 1 + 1 + 1)
    if world == 1 and not args.no_cpu_baseline:
        try:
            line["cpu_baseline"] = cpu_oracle_baseline(S, wl, S.make_camera(W, H))
        except Exception as ex:  # the CPU checker must never take the GPU number down with it
            line["cpu_baseline"] = {"error": str(ex)}
        try:
            line["cpu_baseline"]["torch_cpu"] = torch_cpu_baseline(S)
        except Exception as ex:
            line["cpu_baseline"]["torch_cpu"] = {"error": str(ex)}
    emit(line)
    return 0


if __name__ == "__main__":
    rc = main()
    if torch.distributed.is_available() and torch.distributed.is_initialized():
        torch.distributed.destroy_process_group()
    sys.exit(rc)
