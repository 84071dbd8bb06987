"""All-reduce times of the message sizes the batched-views loop sends (development aid; run under torchrun)."""
import os, sys, datetime, torch, torch.distributed as dist
rank = int(os.environ["RANK"]); world = int(os.environ["WORLD_SIZE"]); local = int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", rank=rank, world_size=world, timeout=datetime.timedelta(seconds=120))
P = 1036800
sizes = {"geometry 11 P": 11 * P, "sh0 all 14 P": 14 * P, "sh3 quarter 12 P": 12 * P, "sh3 half 24 P": 24 * P, "sh3 sh 48 P": 48 * P, "sh3 all 59 P": 59 * P}
if rank == 0:
    print("env", {k: v for k, v in os.environ.items() if k.startswith("NCCL_") and k != "NCCL_DEBUG"}, flush=True)
for name, n in sizes.items():
    x = torch.ones(n, device="cuda")
    for _ in range(5): dist.all_reduce(x)
    torch.cuda.synchronize(); dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(20): dist.all_reduce(x)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 20
    if rank == 0:
        print(f"{name:18s} {4*n/1e6:7.1f} MB  {ms:6.3f} ms  busbw {4*n*2*(world-1)/world/ms/1e6:6.1f} GB/s", flush=True)
n = 48 * P // world * world
x = torch.ones(n, device="cuda"); own = x[rank * (n // world):(rank + 1) * (n // world)]
for name, fn in [("reduce_scatter 48 P", lambda: dist.reduce_scatter_tensor(own, x)), ("all_gather 48 P", lambda: dist.all_gather_into_tensor(x, own))]:
    for _ in range(5): fn()
    torch.cuda.synchronize(); dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(20): fn()
    e1.record(); torch.cuda.synchronize()
    if rank == 0:
        print(f"{name:18s} {4*n/1e6:7.1f} MB  {e0.elapsed_time(e1)/20:6.3f} ms", flush=True)
dist.destroy_process_group()
