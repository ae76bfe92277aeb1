"""Data parallelism: one process per GPU, NCCL over NVLink 5 / NVSwitch (torch.distributed plumbing).

The path shards by batch with three exchange steps, all sums:
  1. gradient all-reduce (asynchronous NCCL all-reduces over the per-parameter gradients, enqueued back to back);
  2. BatchNorm batch statistics, forward [sum x, sum x^2] and backward [sum dy, sum dy*xhat] (fp64, <= 16 KB),
     when sync_bn is on -- an n-rank run then equals the 1-rank run at the same global batch;
  3. the CE / Dice sums of the fused loss (global-batch Dice ratio).
The reference's nn.DataParallel wrap (networks.py:117-121) is degenerate (one id); this replaces it.
"""
from __future__ import annotations

import os
from typing import Iterable, Optional

import torch
import torch.distributed as dist

from . import engine

_GROUP = {"pg": None, "world": 1, "rank": 0}


def _allreduce_sum(t: torch.Tensor) -> None:
    dist.all_reduce(t, op=dist.ReduceOp.SUM, group=_GROUP["pg"])


def init_data_parallel(sync_bn: bool = True, group=None, backend: Optional[str] = None) -> int:
    """Initialise (if needed) torch.distributed from the torchrun env and install the cross-rank reducers.
    Returns the world size."""
    if not dist.is_initialized():
        if "RANK" not in os.environ:
            engine.install_reducer(None, 1, False)
            return 1
        if backend is None:
            backend = "nccl" if torch.cuda.is_available() else "gloo"
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        if backend == "nccl":
            torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", "0")))
        dist.init_process_group(backend=backend)
    _GROUP["pg"] = group
    _GROUP["world"] = dist.get_world_size(group)
    _GROUP["rank"] = dist.get_rank(group)
    engine.install_reducer(_allreduce_sum if _GROUP["world"] > 1 else None, _GROUP["world"], sync_bn)
    return _GROUP["world"]


def shutdown() -> None:
    engine.install_reducer(None, 1, False)
    _GROUP.update(pg=None, world=1, rank=0)


def world_size() -> int:
    return _GROUP["world"]


def rank() -> int:
    return _GROUP["rank"]


def shard_batch(n_global: int, rank_: Optional[int] = None, world: Optional[int] = None):
    """Rank r owns samples [r*B/n, (r+1)*B/n) of the global batch (SURVEY 8e partitioning)."""
    world = world or _GROUP["world"]
    rank_ = _GROUP["rank"] if rank_ is None else rank_
    if n_global % world != 0:
        raise ValueError("global batch %d is not divisible by world size %d" % (n_global, world))
    per = n_global // world
    return rank_ * per, (rank_ + 1) * per


def allreduce_gradients(params: Iterable[torch.Tensor]) -> None:
    """Sum gradients over ranks (averaging is folded into the optimiser's grad_scale = 1/world)."""
    if _GROUP["world"] <= 1:
        return
    grads = [p.grad for p in params if p.grad is not None]
    if not grads:
        return
    # one flat bucket per dtype: a single NCCL all-reduce instead of one latency-bound launch per parameter
    by_dtype = {}
    for g in grads:
        by_dtype.setdefault(g.dtype, []).append(g)
    for gs in by_dtype.values():
        flat = torch.cat([g.reshape(-1) for g in gs])
        dist.all_reduce(flat, op=dist.ReduceOp.SUM, group=_GROUP["pg"])
        torch._foreach_copy_(gs, [c.view_as(g) for c, g in zip(flat.split([g.numel() for g in gs]), gs)])


def broadcast_parameters(module: torch.nn.Module, src: int = 0) -> None:
    if _GROUP["world"] <= 1:
        return
    for t in list(module.parameters()) + list(module.buffers()):
        dist.broadcast(t.data, src=src, group=_GROUP["pg"])
