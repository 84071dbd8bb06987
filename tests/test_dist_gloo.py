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
        # batched global views: the flat gradient reduced in pieces equals the buffer reduced whole, and the pieces tile it
        for M_grad in (1, 16):
            lay = GP.flat_layout(P, M_grad)
            total = lay["_total_floats"][0]
            pieces = GP.comm_pieces(P, M_grad, sh_chunks=3)
            covered = torch.zeros(total, dtype=torch.int32)
            gg = torch.Generator().manual_seed(200 + rank)
            flat = torch.randn(total, generator=gg)
            whole = flat.clone()
            dist.all_reduce(whole)
            works = [dist.all_reduce(flat[off:off + n], async_op=True) for off, n, _, _, _ in pieces]
            for w, (off, n, what, row0, rows) in zip(works, pieces):
                w.wait()
                covered[off:off + n] += 1
                ok &= (what == 1 and off == 0 and n == 11 * P) or (what == 2 and n == 3 * M_grad * rows)
            used = torch.ones(total, dtype=torch.bool)
            used[11 * P:lay["sh"][0]] = False                  # alignment padding between the two blocks
            ok &= bool((covered[used] == 1).all()) and bool((covered[~used] == 0).all())
            ok &= torch.equal(flat[used], whole[used])
            ok &= lay["sh"][0] % 4 == 0 and lay["rotations"][0] == 0
        same = GP.assert_replicas_identical([torch.arange(10, dtype=torch.float32)])
        differ = GP.assert_replicas_identical([torch.full((3,), float(rank))])
        ok &= same and not differ
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
