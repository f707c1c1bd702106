"""Data-parallel plumbing of the learner (host logic; works with any torch.distributed backend).

The path shards per sequence (learner.py:98 is a plain mean over L*B elements, no importance
weights; sampling is with replacement, replay_memory.py:230), so each rank owns a replay shard,
samples its own B sequences and the ONLY exchange is the sum of the flat f32 gradient buffer.
Equal per-rank batch sizes make the mean of the shard means equal to the global-batch mean, so
the rank-summed gradient times 1/world is the gradient of the reference's loss at batch world*B.
"""
from __future__ import annotations

import os

import torch
import torch.distributed as dist


def world_info():
    return (int(os.environ.get("RANK", "0")), int(os.environ.get("LOCAL_RANK", "0")),
            int(os.environ.get("WORLD_SIZE", "1")))


def init_from_env(backend: str | None = None):
    """Create the default process group from torchrun's environment (no-op for world 1)."""
    rank, local, world = world_info()
    if world > 1 and not dist.is_initialized():
        if backend is None:
            backend = "nccl" if torch.cuda.is_available() else "gloo"
        kw = {}
        if backend == "nccl":
            torch.cuda.set_device(local)
            kw["device_id"] = torch.device(f"cuda:{local}")
        dist.init_process_group(backend, **kw)
    return rank, local, world


def active_world(group=None) -> int:
    if dist.is_available() and dist.is_initialized():
        return dist.get_world_size(group)
    return 1


def allreduce_flat_grad(flat_grad: torch.Tensor, group=None) -> float:
    """Sum the flat gradient over ranks in place (one collective per step) and return the
    factor the optimiser kernel must apply (1/world) -- the scale is fused into K10 instead of
    costing another pass over the buffer."""
    world = active_world(group)
    if world == 1:
        return 1.0
    dist.all_reduce(flat_grad, op=dist.ReduceOp.SUM, group=group)
    return 1.0 / world


def broadcast_flat_params(flat: torch.Tensor, src: int = 0, group=None) -> None:
    """Replicas must start from identical weights (the reference has one learner, so this is the
    DP analogue of `target_net.load_state_dict(net.state_dict())`, learner.py:49)."""
    if active_world(group) > 1:
        dist.broadcast(flat, src=src, group=group)
