"""Data parallelism: one process per GPU, NCCL over NVLink 5 / NVSwitch (torch.distributed plumbing).

The path shards by batch with three exchange steps, all sums:
  1. gradient all-reduce (asynchronous NCCL all-reduces over the per-parameter gradients, enqueued back to back);
  2. BatchNorm batch statistics, forward [sum x, sum x^2] and backward [sum dy, sum dy*xhat] (fp64, <= 16 KB),
     when sync_bn is on -- an n-rank run then equals the 1-rank run at the same global batch;
  3. the CE / Dice sums of the fused loss (global-batch Dice ratio).
The reference's nn.DataParallel wrap (networks.py:117-121) is degenerate (one id); this replaces it.

Steps 2 and 3 (36 + 1 latency-bound exchanges on the critical path of a U-Net step) do not go through NCCL on a
single NVLink node: csrc/peer.cu sums them with one kernel per exchange over CUDA-IPC-mapped peer memory (remote
stores + sequence flags, rank-ordered sum -> bit-identical on every rank).  GSB_PEER_ALLREDUCE=0 forces NCCL.
"""
from __future__ import annotations

import os
from typing import Iterable, Optional

import torch
import torch.distributed as dist

from . import engine

_GROUP = {"pg": None, "world": 1, "rank": 0, "peer": None}


def _allreduce_sum(t: torch.Tensor) -> None:
    peer = _GROUP["peer"]
    if peer is not None and peer["world"] == _GROUP["world"] and t.is_cuda and t.dtype == torch.float64 and t.is_contiguous() and t.numel() <= peer["max"]:
        from . import ops
        ops.peer_allreduce_f64(t, peer["regions"], _GROUP["rank"], _GROUP["world"])
        return
    dist.all_reduce(t, op=dist.ReduceOp.SUM, group=_GROUP["pg"])


def _setup_peer_memory(group) -> None:
    """Create this rank's region, exchange CUDA IPC handles, map the peers (same node, <= 8 ranks, NCCL backend)."""
    if _GROUP["peer"] is not None or os.environ.get("GSB_PEER_ALLREDUCE", "1") == "0":
        return
    world, rank_ = _GROUP["world"], _GROUP["rank"]
    local_world = int(os.environ.get("LOCAL_WORLD_SIZE", str(world)))
    if world < 2 or world > 8 or local_world != world or dist.get_backend(group) != "nccl":
        return
    from . import ops
    region, handle = ops.peer_create()
    handles = [None] * world
    dist.all_gather_object(handles, handle, group=group)
    regions = [region if r == rank_ else ops.peer_open(handles[r]) for r in range(world)]
    dist.barrier(group=group)    # every region is zeroed and mapped before the first exchange
    _GROUP["peer"] = {"regions": regions, "max": ops.peer_max_elems(), "own": region, "world": world}


class _Bf16Bucket:
    """An fp32 gradient bucket whose all-reduce travels as bf16 (half the NVLink bytes): cast -> asynchronous NCCL
    all-reduce of the bf16 copy -> ``wait()`` casts the sum back over the fp32 bucket.  The fp32 tensor stays the
    parameter's ``.grad`` storage."""

    def __init__(self, t: torch.Tensor):
        from . import ops
        self.t = t
        self.buf = torch.empty(t.numel(), device=t.device, dtype=torch.bfloat16)
        ops.cast_flat(t, self.buf)
        self.work = dist.all_reduce(self.buf, op=dist.ReduceOp.SUM, group=_GROUP["pg"], async_op=True)

    def wait(self) -> None:
        from . import ops
        self.work.wait()
        ops.cast_flat(self.buf, self.t)


def _allreduce_async(t: torch.Tensor):
    """Asynchronous in-place sum of a flat fp32 gradient bucket over the ranks -> handle with .wait().
    GSB_GRAD_COMM_DTYPE=bf16 (set by init_data_parallel(grad_comm_dtype=...)) sends the bucket as bf16."""
    if (_GROUP.get("grad_dtype") == torch.bfloat16 and t.is_cuda and t.dtype == torch.float32 and t.is_contiguous()
            and t.data_ptr() % 16 == 0 and t.numel() >= 4096):
        return _Bf16Bucket(t)
    return dist.all_reduce(t, op=dist.ReduceOp.SUM, group=_GROUP["pg"], async_op=True)


def init_data_parallel(sync_bn: bool = True, group=None, backend: Optional[str] = None,
                       overlap_grad_allreduce: bool = False, grad_comm_dtype: Optional[str] = None) -> int:
    """Initialise (if needed) torch.distributed from the torchrun env and install the cross-rank reducers.
    Returns the world size.

    overlap_grad_allreduce: the U-Net program all-reduces its gradients INSIDE backward -- the decoder + bottleneck
    bucket (72 % of the bytes) as soon as it is final, on the collective's stream, under the encoder backward; the
    rest at the end -- and marks the parameters so that ``allreduce_gradients`` skips them.  Use it with
    ``loss.backward()`` training loops (``torch.autograd.grad`` callers keep it off).

    grad_comm_dtype: 'fp32' (default; an N-rank run then equals the 1-rank run up to fp32 summation order) or 'bf16'
    (gradient buckets are cast to bf16 for the all-reduce: half the bytes, 2^-9 relative rounding per addend); the
    environment variable GSB_GRAD_COMM_DTYPE overrides."""
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
    gd = os.environ.get("GSB_GRAD_COMM_DTYPE", grad_comm_dtype or "fp32").lower()
    if gd not in ("fp32", "bf16"):
        raise ValueError("grad_comm_dtype must be 'fp32' or 'bf16'")
    _GROUP["grad_dtype"] = torch.bfloat16 if gd == "bf16" else torch.float32
    _GROUP["pg"] = group
    _GROUP["world"] = dist.get_world_size(group)
    _GROUP["rank"] = dist.get_rank(group)
    if backend != "gloo" and torch.cuda.is_available():
        _setup_peer_memory(group)
    peer = _GROUP["peer"]
    peer_arg = None
    if peer is not None and peer["world"] == _GROUP["world"] and _GROUP["world"] > 1:
        peer_arg = {"regions": peer["regions"], "rank": _GROUP["rank"], "world": _GROUP["world"]}
    multi = _GROUP["world"] > 1
    engine.install_reducer(_allreduce_sum if multi else None, _GROUP["world"], sync_bn, peer_arg,
                           grad_fn=_allreduce_async if (multi and overlap_grad_allreduce) else None,
                           sm_reserve=int(os.environ.get("GSB_OVERLAP_SM_RESERVE", "16")) if multi and overlap_grad_allreduce else 0,
                           rank=_GROUP["rank"])
    return _GROUP["world"]


def check_peer_health() -> None:
    """Raise if one of this rank's peer-memory exchanges (sync-BN statistics, CE / Dice sums) gave up waiting for a peer
    (GSB_PEER_TIMEOUT_S, default 30 minutes).  The kernels never trap: they record the call, continue with the local value
    and return, so the CUDA context stays usable and the failure surfaces here as an ordinary exception.  Synchronises the
    device: call it at step / epoch boundaries (``evaluate`` and ``shutdown`` do).  Every rank must issue the same sequence
    of train-mode BatchNorm layers and ``seg_loss`` calls; a rank that skips one (a validation loss only rank 0 computes)
    leaves the others waiting."""
    peer = _GROUP["peer"]
    if peer is None:
        return
    from . import ops
    seq = ops.peer_error(peer["own"])
    if seq:
        raise RuntimeError("genseg_b200: peer-memory exchange #%d on rank %d timed out waiting for a peer "
                           "(GSB_PEER_TIMEOUT_S); the ranks did not issue the same sequence of sync-BN / seg_loss calls, "
                           "or a peer died" % (seq, _GROUP["rank"]))


def shutdown() -> None:
    if _GROUP["peer"] is not None and _GROUP["world"] > 1:
        check_peer_health()
    engine.install_reducer(None, 1, False)
    _GROUP.update(pg=None, world=1, rank=0)   # peer regions stay mapped: a later init_data_parallel reuses them


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
    """Sum gradients over ranks.  Whether the optimiser must then divide depends on the loss: ``seg_loss`` under
    init_data_parallel is a mean over the GLOBAL batch (its CE count and Dice sums are all-reduced), so the sum of the
    ranks' gradients already is the full-batch gradient (``grad_scale = 1``); per-rank-mean losses (GANLoss, l1_loss)
    need ``optimizer.grad_scale = 1 / world``."""
    if _GROUP["world"] <= 1:
        return
    grads = []
    for p in params:
        if getattr(p, "_gsb_grad_reduced", False):   # already summed inside backward (overlap_grad_allreduce)
            p._gsb_grad_reduced = False
            continue
        if p.grad is not None:
            grads.append(p.grad)
    if not grads:
        return
    # Large contiguous gradients (the generator's 64 MB up-cell kernels) are reduced IN PLACE, back to back and
    # asynchronously: no concatenated copy of a 1 GB gradient set.  The small ones (BatchNorm, biases, thin layers) share
    # one flat bucket per dtype: a single all-reduce instead of one latency-bound launch per parameter.
    big, small = [], {}
    for g in grads:
        if g.is_contiguous() and g.numel() >= (1 << 20):
            big.append(g)
        else:
            small.setdefault(g.dtype, []).append(g)
    handles = [_allreduce_async(g.view(-1)) for g in big]
    for gs in small.values():
        flat = torch.cat([g.reshape(-1) for g in gs])
        dist.all_reduce(flat, op=dist.ReduceOp.SUM, group=_GROUP["pg"])
        torch._foreach_copy_(gs, [c.view_as(g) for c, g in zip(flat.split([g.numel() for g in gs]), gs)])
    for h in handles:
        h.wait()


def mean_over_ranks(total, count: int) -> torch.Tensor:
    """Evaluation is "replicas only" (SURVEY 8e): every rank scores its own batches and the ranks only exchange two numbers,
    the sum of the per-batch Dice scores and the number of batches.  Returns sum(total) / sum(count) over the ranks."""
    t = torch.as_tensor(total, dtype=torch.float64).detach().reshape(())
    if _GROUP["world"] <= 1:
        return (t / max(int(count), 1)).float()
    if dist.get_backend(_GROUP["pg"]) == "nccl" and not t.is_cuda:
        # a rank that scored no batch holds the Python number 0: the collective still needs a tensor on this rank's GPU
        t = t.to(torch.device("cuda", torch.cuda.current_device()))
    pair = torch.stack([t, torch.tensor(float(count), dtype=torch.float64, device=t.device)])
    dist.all_reduce(pair, op=dist.ReduceOp.SUM, group=_GROUP["pg"])
    return (pair[0] / pair[1].clamp_min(1.0)).float()


def broadcast_parameters(module: torch.nn.Module, src: int = 0) -> None:
    if _GROUP["world"] <= 1:
        return
    for t in list(module.parameters()) + list(module.buffers()):
        dist.broadcast(t.data, src=src, group=_GROUP["pg"])
