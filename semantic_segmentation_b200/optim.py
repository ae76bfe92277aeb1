"""Fused optimisers on the sm_100a kernels, torch.optim-shaped.

RMSprop mirrors torch.optim.RMSprop as used at running_files/train_end2end_jsrt.py:71-72
(lr, weight_decay=1e-8, momentum=0.9, alpha=0.99, eps=1e-8); Adam mirrors torch.optim.Adam as used at
models_pix2pix/pix2pix_model.py:75-78 (betas=(0.5, 0.999)).  ``grad_scale`` folds the data-parallel
1/world averaging into the update.
"""
from __future__ import annotations

from typing import Iterable, List

import torch

from . import ops


class _Base:
    def __init__(self, params: Iterable[torch.Tensor]):
        self.params: List[torch.Tensor] = [p for p in params]
        self.state = {}
        self.grad_scale = 1.0

    def zero_grad(self, set_to_none: bool = True) -> None:
        for p in self.params:
            if set_to_none:
                p.grad = None
            elif p.grad is not None:
                p.grad.zero_()


class RMSprop(_Base):
    def __init__(self, params, lr=1e-2, alpha=0.99, eps=1e-8, weight_decay=0.0, momentum=0.0):
        super().__init__(params)
        self.lr, self.alpha, self.eps, self.weight_decay, self.momentum = lr, alpha, eps, weight_decay, momentum

    @torch.no_grad()
    def step(self) -> None:
        ps, gs, sqs, bufs = [], [], [], []
        for p in self.params:
            if p.grad is None:
                continue
            st = self.state.get(p)
            if st is None:
                st = self.state[p] = (torch.zeros_like(p), torch.zeros_like(p) if self.momentum > 0 else None)
            ps.append(p.data)
            gs.append(p.grad if p.grad.is_contiguous() else p.grad.contiguous())
            sqs.append(st[0])
            bufs.append(st[1])
        # one launch over every parameter tensor (multi-tensor kernel)
        ops.rmsprop_step_multi(ps, gs, sqs, bufs, self.lr, self.alpha, self.eps, self.weight_decay, self.momentum,
                               self.grad_scale)


class Adam(_Base):
    def __init__(self, params, lr=1e-3, betas=(0.9, 0.999), eps=1e-8, weight_decay=0.0):
        super().__init__(params)
        self.lr, self.betas, self.eps, self.weight_decay = lr, betas, eps, weight_decay
        self.t = 0

    @torch.no_grad()
    def step(self) -> None:
        self.t += 1
        ps, gs, ms, vs = [], [], [], []
        for p in self.params:
            if p.grad is None:
                continue
            st = self.state.get(p)
            if st is None:
                st = self.state[p] = (torch.zeros_like(p), torch.zeros_like(p))
            ps.append(p.data)
            gs.append(p.grad if p.grad.is_contiguous() else p.grad.contiguous())
            ms.append(st[0])
            vs.append(st[1])
        ops.adam_step_multi(ps, gs, ms, vs, self.lr, self.betas[0], self.betas[1], self.eps, self.weight_decay, self.t,
                            self.grad_scale)
