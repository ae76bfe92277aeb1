"""Fused optimisers on the sm_100a kernels: ``torch.optim.Optimizer`` subclasses, so everything the
reference hangs on its optimisers keeps working -- ``param_groups`` (``optim.lr_scheduler.ReduceLROnPlateau``,
train_end2end_jsrt.py:73), ``state_dict()`` / ``load_state_dict()`` (checkpoints, Betty's roll_back snapshot),
``zero_grad`` -- while ``step()`` is ONE multi-tensor launch per parameter group.

RMSprop mirrors torch.optim.RMSprop as used at running_files/train_end2end_jsrt.py:71-72
(lr, weight_decay=1e-8, momentum=0.9, alpha=0.99, eps=1e-8); Adam mirrors torch.optim.Adam as used at
models_pix2pix/pix2pix_model.py:75-78 (betas=(0.5, 0.999)).  The per-parameter state uses torch's key names
(``square_avg`` / ``momentum_buffer``, ``exp_avg`` / ``exp_avg_sq``, ``step``), so an optimiser checkpoint written by the
reference loads here and vice versa (``step`` is kept as a Python number; torch's ``__setstate__`` converts it).
``grad_scale`` folds the data-parallel 1/world averaging into the update.  There is no CPU path: ``step()`` on CPU
parameters raises RuntimeError before touching any state.
"""
from __future__ import annotations

import torch
from torch.optim import Optimizer

from . import ops


def _grad(p: torch.Tensor) -> torch.Tensor:
    return p.grad if p.grad.is_contiguous() else p.grad.contiguous()


def _require_cuda(param_groups) -> None:
    for group in param_groups:
        for p in group["params"]:
            if not p.is_cuda:
                raise RuntimeError("genseg_b200: the fused optimisers need CUDA parameters; there is no CPU fallback")


class RMSprop(Optimizer):
    def __init__(self, params, lr=1e-2, alpha=0.99, eps=1e-8, weight_decay=0.0, momentum=0.0, centered=False, foreach=None):
        if centered:
            raise NotImplementedError("genseg_b200: centered RMSprop is not on the B200 path (the reference does not use it)")
        if lr < 0.0 or eps < 0.0 or alpha < 0.0 or momentum < 0.0 or weight_decay < 0.0:
            raise ValueError("invalid RMSprop hyper-parameter")
        super().__init__(params, dict(lr=lr, alpha=alpha, eps=eps, weight_decay=weight_decay, momentum=momentum))
        self.grad_scale = 1.0

    @torch.no_grad()
    def step(self, closure=None):
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        _require_cuda(self.param_groups)
        for group in self.param_groups:
            momentum = group["momentum"]
            ps, gs, sqs, bufs = [], [], [], []
            for p in group["params"]:
                if p.grad is None:
                    continue
                st = self.state[p]
                if "square_avg" not in st:
                    st["step"] = 0
                    st["square_avg"] = torch.zeros_like(p)
                    if momentum > 0:
                        st["momentum_buffer"] = torch.zeros_like(p)
                elif momentum > 0 and "momentum_buffer" not in st:
                    st["momentum_buffer"] = torch.zeros_like(p)
                st["step"] = st["step"] + 1
                ps.append(p.data)
                gs.append(_grad(p))
                sqs.append(st["square_avg"])
                bufs.append(st.get("momentum_buffer"))
            # one launch over every parameter tensor of the group (multi-tensor kernel)
            ops.rmsprop_step_multi(ps, gs, sqs, bufs, float(group["lr"]), group["alpha"], group["eps"], group["weight_decay"],
                                   momentum, self.grad_scale)
        return loss


class Adam(Optimizer):
    def __init__(self, params, lr=1e-3, betas=(0.9, 0.999), eps=1e-8, weight_decay=0.0, amsgrad=False, foreach=None,
                 capturable=False):
        if amsgrad:
            raise NotImplementedError("genseg_b200: amsgrad is not on the B200 path (the reference does not use it)")
        if lr < 0.0 or eps < 0.0 or not 0.0 <= betas[0] < 1.0 or not 0.0 <= betas[1] < 1.0 or weight_decay < 0.0:
            raise ValueError("invalid Adam hyper-parameter")
        super().__init__(params, dict(lr=lr, betas=tuple(betas), eps=eps, weight_decay=weight_decay))
        self.grad_scale = 1.0
        # capturable (torch.optim.Adam's flag): the step count lives in device memory (one int64 scalar per parameter
        # group, state["step"] of every parameter refers to it) and the kernel forms the bias corrections from it, so a
        # training step captured into a CUDA graph advances at every replay
        self.capturable = bool(capturable)
        self._step_dev = {}

    def _step_capturable(self, gi: int, group) -> None:
        params = [p for p in group["params"] if p.grad is not None]
        if not params:
            return
        ctr = self._step_dev.get(gi)
        if ctr is None:
            t0 = max([int(self.state[p].get("step", 0)) for p in params if "exp_avg" in self.state[p]] or [0])
            ctr = self._step_dev[gi] = torch.full((), t0, device=params[0].device, dtype=torch.int64)
        ps, gs, ms, vs = [], [], [], []
        for p in params:
            st = self.state[p]
            if "exp_avg" not in st:
                st["exp_avg"] = torch.zeros_like(p)
                st["exp_avg_sq"] = torch.zeros_like(p)
            st["step"] = ctr
            ps.append(p.data)
            gs.append(_grad(p))
            ms.append(st["exp_avg"])
            vs.append(st["exp_avg_sq"])
        ops.counter_add(ctr, 1)
        ops.adam_step_multi_dev(ps, gs, ms, vs, float(group["lr"]), group["betas"][0], group["betas"][1], group["eps"],
                                group["weight_decay"], ctr, self.grad_scale)

    @torch.no_grad()
    def step(self, closure=None):
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        _require_cuda(self.param_groups)
        if self.capturable:
            for gi, group in enumerate(self.param_groups):
                self._step_capturable(gi, group)
            return loss
        for group in self.param_groups:
            by_step = {}                                   # bias correction depends on the step count: one launch per count
            for p in group["params"]:
                if p.grad is None:
                    continue
                st = self.state[p]
                if "exp_avg" not in st:
                    st["step"] = 0
                    st["exp_avg"] = torch.zeros_like(p)
                    st["exp_avg_sq"] = torch.zeros_like(p)
                t = int(st["step"]) + 1
                st["step"] = t
                ps, gs, ms, vs = by_step.setdefault(t, ([], [], [], []))
                ps.append(p.data)
                gs.append(_grad(p))
                ms.append(st["exp_avg"])
                vs.append(st["exp_avg_sq"])
            for t, (ps, gs, ms, vs) in by_step.items():
                ops.adam_step_multi(ps, gs, ms, vs, float(group["lr"]), group["betas"][0], group["betas"][1], group["eps"],
                                    group["weight_decay"], t, self.grad_scale)
        return loss
