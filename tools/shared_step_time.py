"""Time the shared optimiser step (gsr_gaussian_adam_step_shared over NVSwitch multicast) on its own, under torchrun:
two barriers alone, and barrier + step + barrier at SH degree 0 and 3 (development aid; device events, max over ranks)."""
import json, os, sys
import torch
import torch.distributed as dist
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "cf-3dgs_b200"))


def main():
    dist.init_process_group("nccl")
    rank, world = dist.get_rank(), dist.get_world_size()
    dev = torch.device("cuda", int(os.environ.get("LOCAL_RANK", 0)))
    torch.cuda.set_device(dev)
    import gsr_model, gsr_parallel
    P, M = int(os.environ.get("GSR_P", 1036800)), 16
    g = torch.Generator().manual_seed(1)
    mk = lambda *s: torch.randn(*s, generator=g).to(dev)
    lrs = dict(xyz=1.6e-4, f_dc=2.5e-3, f_rest=1.25e-4, opacity=0.05, scaling=5e-3, rotation=1e-3)
    model = gsr_model.FusedGaussianModel(mk(P, 3), 0.3 * mk(P, M, 3), mk(P, 1), mk(P, 3) * 0.3 - 3, mk(P, 4), lrs, device_step=True)
    bvs = gsr_parallel.BatchedViewsStep(model, multicast=True)
    mc = bvs._mc
    radii = torch.zeros(P, dtype=torch.int32, device=dev)
    mc["buf"][mc["per"] * mc["Pp"]:].normal_(generator=None)       # some gradient

    def timed(fn, n=50, warm=5):
        for _ in range(warm):
            fn()
        dist.barrier(); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda._sleep(int(60e-3 * 1.9e9))      # the host enqueues all n calls behind this: device time, not Python's
        e0.record()
        for _ in range(n):
            fn()
        e1.record(); torch.cuda.synchronize()
        t = torch.tensor([e0.elapsed_time(e1) / n], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t)

    out = {"n_gpus": world, "P": P}
    out["two_barriers_ms"] = timed(lambda: (mc["hdl"].barrier(channel=0), mc["hdl"].barrier(channel=0)))
    for deg in (0, 3):
        Mg = (deg + 1) ** 2
        model.sh_ever = Mg
        model.shs(deg)             # degree 0: the dense coefficient-0 plane (in the symmetric buffer) is what gets stepped
        out[f"shared_update_sh{deg}_ms"] = timed(lambda: bvs._shared_update(Mg, radii))
    if os.environ.get("GSR_PIECES", "1") == "0":
        if rank == 0:
            print(json.dumps(out), flush=True)
        dist.destroy_process_group()
        return
    # the pieces: the per-Gaussian block alone (no SH rows), and the same kernels on plain addresses (no NVLink at all)
    import ctypes
    from diff_gaussian_rasterization._cabi import check, lib
    lr = (ctypes.c_float * 6)(1.6e-4, 2.5e-3, 1.25e-4, 0.05, 5e-3, 1e-3)
    Pp, per, buf = mc["Pp"], mc["per"], mc["buf"]
    order = ("xyz", "features", "opacity_raw", "log_scales", "rot_raw")
    local = [buf.data_ptr() + 4 * mc["par"][k][0] for k in order]
    mcp = [mc["mc0"] + 4 * mc["par"][k][0] for k in order]
    lgrad = [buf.data_ptr() + 4 * (per * Pp + mc["par"][k][0]) for k in order]
    mgrad = [mc["mc0"] + 4 * (per * Pp + mc["par"][k][0]) for k in order]

    def raw(multicast, M_grad, M_upd):
        s = torch.cuda.current_stream().cuda_stream
        mc["hdl"].barrier(channel=0)
        check(lib.gsr_gaussian_adam_step_shared(Pp, M, M_grad, M_upd, rank * mc["rows_per"], mc["rows_per"], int(multicast), 1,
                                                model.step_state.data_ptr(), 0.9, 0.999, 1e-15, lr, *local,
                                                *(mcp if multicast else local), *(mgrad if multicast else lgrad),
                                                model.state.data_ptr(), s))
        mc["hdl"].barrier(channel=0)
    out["geometry_only_multicast_ms"] = timed(lambda: raw(True, 1, 0))
    out["geometry_only_plain_ms"] = timed(lambda: raw(False, 1, 0))
    out["sh0_multicast_ms"] = timed(lambda: raw(True, 1, 1))
    out["sh0_plain_ms"] = timed(lambda: raw(False, 1, 1))
    out["sh3_multicast_ms"] = timed(lambda: raw(True, 16, 16))
    out["sh3_plain_ms"] = timed(lambda: raw(False, 16, 16))
    ok = gsr_parallel.assert_replicas_identical([model.xyz, model.features, model.opacity_raw, model.log_scales, model.rot_raw])
    out["replicas_bit_identical"] = ok
    if rank == 0:
        print(json.dumps(out), flush=True)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
