"""Multi-GPU plumbing for the two ways the CF-3DGS optimisation shards.

One process per GPU, torch.distributed (NCCL on the B200 box, gloo in CPU tests).  The rasterizer for one view is
never split ("replicas only" inside a view); across the optimisation:
  (i)  independent frame pairs: pair i -> rank i mod world, no data-path collective while fitting, then ONE
       all-gather of the 4x4 relative poses and a sequential prefix product
       (trainer/cf3dgs_trainer.py:225-297: pose = rel_pose @ get_RT(fidx-1));
  (ii) batched global views: every rank renders its own view of a replicated model, then one all-reduce(sum) of the
       flat global-Gaussian gradient (xyz 3 + f_dc 3 + f_rest 45 + opacity 1 + scaling 3 + rotation 4 = 59 floats
       per Gaussian, scene/gaussian_model_cf.py:254-266).
"""
from __future__ import annotations

from typing import Dict, List, Sequence

import torch
import torch.distributed as dist

GRAD_ORDER = ("means3D", "shs", "opacities", "scales", "rotations")


def shard_pairs(n_pairs: int, rank: int, world: int) -> List[int]:
    """Frame pairs (i-1, i) owned by this rank: round-robin so early and late frames mix."""
    return [i for i in range(n_pairs) if i % world == rank]


def flat_size(grads: Dict[str, torch.Tensor]) -> int:
    return sum(grads[k].numel() for k in GRAD_ORDER if k in grads and grads[k] is not None)


def flatten_into(grads: Dict[str, torch.Tensor], flat: torch.Tensor) -> torch.Tensor:
    off = 0
    for k in GRAD_ORDER:
        g = grads.get(k)
        if g is None:
            continue
        n = g.numel()
        flat[off:off + n].copy_(g.reshape(-1))
        off += n
    return flat[:off]


def unflatten_from(flat: torch.Tensor, like: Dict[str, torch.Tensor]) -> Dict[str, torch.Tensor]:
    out, off = {}, 0
    for k in GRAD_ORDER:
        g = like.get(k)
        if g is None:
            continue
        n = g.numel()
        out[k] = flat[off:off + n].view_as(g)
        off += n
    return out


def allreduce_gradients(grads: Dict[str, torch.Tensor], flat: torch.Tensor, average: bool = False):
    """Sum (or average) the per-rank gradients of a replicated Gaussian model with one collective."""
    used = flatten_into(grads, flat)
    if dist.is_initialized() and dist.get_world_size() > 1:
        dist.all_reduce(used)
        if average:
            used.div_(dist.get_world_size())
    return unflatten_from(used, grads)


def allreduce_densify_stats(grad_accum: torch.Tensor, denom: torch.Tensor, max_radii2D: torch.Tensor):
    """Densification statistics are (sum, sum, max) across ranks (scene/gaussian_model_cf.py:716-719,
    trainer/cf3dgs_trainer.py:147-152)."""
    if dist.is_initialized() and dist.get_world_size() > 1:
        dist.all_reduce(grad_accum)
        dist.all_reduce(denom)
        dist.all_reduce(max_radii2D, op=dist.ReduceOp.MAX)
    return grad_accum, denom, max_radii2D


def gather_relative_poses(local: Dict[int, torch.Tensor], n_pairs: int) -> List[torch.Tensor]:
    """All-gather the 4x4 relative poses each rank fitted; returns them indexed by pair."""
    world = dist.get_world_size() if dist.is_initialized() else 1
    dev = next(iter(local.values())).device if local else torch.device("cpu")
    buf = torch.zeros(n_pairs, 4, 4, device=dev)
    for i, T in local.items():
        buf[i] = T
    if world > 1:
        dist.all_reduce(buf)   # disjoint ownership: a sum is a gather
    return [buf[i] for i in range(n_pairs)]


def compose_poses(rel: Sequence[torch.Tensor], first: torch.Tensor | None = None) -> List[torch.Tensor]:
    """Absolute poses by the reference's recurrence pose_i = rel_i @ pose_{i-1}
    (trainer/cf3dgs_trainer.py:294-297)."""
    cur = torch.eye(4, device=rel[0].device, dtype=rel[0].dtype) if first is None else first
    out = [cur]
    for T in rel:
        cur = T @ cur
        out.append(cur)
    return out
