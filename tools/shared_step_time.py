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
        out[f"shared_update_sh{deg}_ms"] = timed(lambda: bvs._shared_update(Mg, radii))
    ok = gsr_parallel.assert_replicas_identical([model.xyz, model.features, model.opacity_raw, model.log_scales, model.rot_raw])
    out["replicas_bit_identical"] = ok
    if rank == 0:
        print(json.dumps(out), flush=True)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
