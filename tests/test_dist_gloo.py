"""CPU, world_size 2, gloo: host-side logic of the N>1 paths."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import gsr_parallel as GP


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close()
    return p


def _worker(rank, world, port, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        P = 50
        g = torch.Generator().manual_seed(100 + rank)
        grads = {"means3D": torch.randn(P, 3, generator=g), "shs": torch.randn(P, 16, 3, generator=g),
                 "opacities": torch.randn(P, 1, generator=g), "scales": torch.randn(P, 3, generator=g),
                 "rotations": torch.randn(P, 4, generator=g)}
        assert GP.flat_size(grads) == P * 59
        flat = torch.empty(P * 59)
        red = GP.allreduce_gradients({k: v.clone() for k, v in grads.items()}, flat)
        # expected: sum over both ranks' seeded gradients
        exp = {}
        for r in range(world):
            gg = torch.Generator().manual_seed(100 + r)
            one = {"means3D": torch.randn(P, 3, generator=gg), "shs": torch.randn(P, 16, 3, generator=gg),
                   "opacities": torch.randn(P, 1, generator=gg), "scales": torch.randn(P, 3, generator=gg),
                   "rotations": torch.randn(P, 4, generator=gg)}
            for k, v in one.items():
                exp[k] = exp.get(k, 0) + v
        ok = all(torch.allclose(red[k], exp[k], atol=1e-6) for k in exp)
        # densification statistics: sum, sum, max
        a, d, m = GP.allreduce_densify_stats(torch.full((4,), float(rank + 1)), torch.ones(4), torch.tensor([float(rank), 5.0 - rank, 1.0, 2.0]))
        ok &= torch.equal(a, torch.full((4,), 3.0)) and torch.equal(d, torch.full((4,), 2.0)) and torch.equal(m, torch.tensor([1.0, 5.0, 1.0, 2.0]))
        # frame pairs: disjoint ownership, gather, prefix product
        n_pairs = 5
        mine = GP.shard_pairs(n_pairs, rank, world)
        local = {}
        for i in mine:
            T = torch.eye(4); T[0, 3] = float(i + 1)
            local[i] = T
        rel = GP.gather_relative_poses(local, n_pairs)
        poses = GP.compose_poses(rel)
        ok &= len(poses) == n_pairs + 1 and abs(float(poses[-1][0, 3]) - 15.0) < 1e-6
        q.put((rank, bool(ok), mine))
    finally:
        dist.destroy_process_group()


def test_two_rank_gradient_allreduce_and_pose_gather():
    world, port = 2, _free_port()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
    assert all(ok for _, ok, _ in res)
    owned = sorted(i for _, _, mine in res for i in mine)
    assert owned == list(range(5))


def test_shard_pairs_partition():
    for world in (1, 2, 4, 8):
        allp = sorted(i for r in range(world) for i in GP.shard_pairs(119, r, world))
        assert allp == list(range(119))
        sizes = [len(GP.shard_pairs(119, r, world)) for r in range(world)]
        assert max(sizes) - min(sizes) <= 1
