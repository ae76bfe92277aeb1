"""Hyper-gradient along a dependency path (PARITY UNPINNED, see the package docstring).

For an upper level with variables lam and loss L_up(w*(lam), lam) over a lower level w*(lam) = argmin_w L_low(w, lam):

    dL_up/dlam = dL_up/dlam|direct + (dw*/dlam)^T v,        v = dL_up/dw

DARTS (Liu et al. 2019, eq. 7-8) approximates w* by one gradient step, so dw*/dlam ~ -xi * d2 L_low / (dlam dw), and the
Hessian-vector product by a central finite difference of lower-level gradients at w +- eps * v:

    (dw*/dlam)^T v  ~  -( grad_lam L_low(w + eps v, lam) - grad_lam L_low(w - eps v, lam) ) / (2 eps),   eps = alpha / ||v||

Betty's ``darts`` drops the step size xi (it is absorbed in the upper optimiser's learning rate); so does this.
Each evaluation of L_low re-runs the lower problem's ``training_step`` on ITS current batch (``curr.cur_batch``), with the
parameters perturbed in place through ``p.data`` -- which is why the CUDA modules re-pack their weights every forward
(SURVEY fact 10).
"""
from __future__ import annotations

from typing import List, Optional, Sequence

import torch
import torch.distributed as dist


def allreduce_mean(tensors: Sequence[Optional[torch.Tensor]]) -> None:
    """In-place average over the ranks of every non-None tensor (one flat bucket per dtype)."""
    world = dist.get_world_size()
    by_dtype = {}
    for t in tensors:
        if t is not None:
            by_dtype.setdefault(t.dtype, []).append(t)
    for ts in by_dtype.values():
        flat = torch.cat([t.reshape(-1) for t in ts])
        dist.all_reduce(flat, op=dist.ReduceOp.SUM)
        flat.div_(world)
        for t, c in zip(ts, flat.split([t.numel() for t in ts])):
            t.copy_(c.view_as(t))


def _distributed(sync: bool) -> bool:
    return bool(sync and dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1)


def _loss_of(curr):
    out = curr.training_step_exec(curr.cur_batch)
    if isinstance(out, dict):
        return out["loss"]
    if isinstance(out, (tuple, list)):
        return out[0]
    return out


def darts(vector: List[Optional[torch.Tensor]], curr, prev, sync: bool = False) -> List[Optional[torch.Tensor]]:
    """Implicit gradient w.r.t. ``prev.trainable_parameters()`` of <w*(prev), vector>, w* being ``curr``'s variables.

    ``vector`` is aligned with ``curr.trainable_parameters()`` (None = zero).  Under data parallelism the direction is
    averaged over the ranks first, so that every rank perturbs its replica identically."""
    params = list(curr.trainable_parameters())
    targets = list(prev.trainable_parameters())
    if _distributed(sync):
        vector = [torch.zeros_like(p) if v is None else v.detach().clone() for p, v in zip(params, vector)]
        allreduce_mean(vector)
    live = [(p, v.detach()) for p, v in zip(params, vector) if v is not None]
    if not live or not targets:
        return [None] * len(targets)
    norm = torch.linalg.vector_norm(torch.stack([torch.linalg.vector_norm(v.float()) for _, v in live])).item()
    if not norm > 0.0:       # the upper loss does not reach this level: no implicit term (also catches NaN)
        return [None] * len(targets)
    eps = curr.config.darts_alpha / norm

    def perturbed_grads(scale):
        with torch.no_grad():
            for p, v in live:
                p.data.add_(v.to(p.dtype), alpha=scale)
        loss = _loss_of(curr)
        return torch.autograd.grad(loss, targets, allow_unused=True)

    grad_p = perturbed_grads(eps)             # w + eps v
    grad_n = perturbed_grads(-2.0 * eps)      # w - eps v
    with torch.no_grad():
        for p, v in live:
            p.data.add_(v.to(p.dtype), alpha=eps)   # back to w
    out: List[Optional[torch.Tensor]] = []
    for gp, gn in zip(grad_p, grad_n):
        if gp is None and gn is None:
            out.append(None)
        else:
            gp = torch.zeros_like(gn) if gp is None else gp
            gn = torch.zeros_like(gp) if gn is None else gn
            out.append((gn - gp).div_(2.0 * eps))
    return out


_GRAD_FN = {"darts": darts}


def get_grads(loss: torch.Tensor, path, retain_graph: bool = False, sync: bool = False) -> List[Optional[torch.Tensor]]:
    """Chain the lower levels of ``path`` = [upper, lower_k, ..., lower_1, upper]: the first vector is dL_up / d lower_k,
    every hop maps a vector on path[i] to one on path[i + 1]; the result is aligned with upper.trainable_parameters()."""
    if len(path) < 3 or path[0] is not path[-1]:
        raise ValueError("a hyper-gradient path is [upper, lower..., upper]")
    first = list(path[1].trainable_parameters())
    if not first:
        return [None] * len(list(path[-1].trainable_parameters()))
    vector = list(torch.autograd.grad(loss, first, retain_graph=retain_graph, allow_unused=True))
    upper = list(path[-1].trainable_parameters())
    for i in range(1, len(path) - 1):
        fn = _GRAD_FN[path[i].config.type]
        vector = fn(vector, path[i], path[i + 1], sync=sync)
        if all(v is None for v in vector):
            return [None] * len(upper)
    return vector
