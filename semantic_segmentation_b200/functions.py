"""torch.autograd glue: ONE autograd node per network call.

``ProgramFunction.apply(program, module_buffers, training, extra, n_in, grad_mode, *inputs, *params)`` runs the program's explicit
forward on the CUDA kernels and keeps its tape; backward runs the explicit backward.  It supports
loss.backward(), torch.autograd.grad(..., allow_unused=True), retain_graph=True (the tape is never
mutated by backward), torch.no_grad() (nothing saved) and requires_grad_(False) parameters (their
wgrad kernels are skipped) -- the autograd contract of SURVEY.md section 8b.
"""
from __future__ import annotations

from typing import Dict, List

import torch

from ._lib import require_cuda


class ProgramFunction(torch.autograd.Function):
    @staticmethod
    def forward(ctx, program, buffers: Dict[str, torch.Tensor], training: bool, extra, n_in: int, grad_mode: bool, *tensors):
        xs, params = tensors[:n_in], tensors[n_in:]
        for x in xs:
            require_cuda(x, "input")
        names: List[str] = program.param_names
        if len(params) != len(names):
            raise RuntimeError("genseg_b200: expected %d parameters, got %d" % (len(names), len(params)))
        P = {k: v.detach() for k, v in zip(names, params)}
        for k, v in P.items():
            require_cuda(v, k)
        # ctx.needs_input_grad reflects requires_grad only -- it stays True under torch.no_grad() -- so the caller's grad mode
        # travels as an argument: without it nothing is saved, no backward weight layout is packed and, in eval mode, the
        # BatchNorm folds into the convolution epilogues (engine.ConvBNAct.forward)
        need_any = bool(grad_mode) and any(ctx.needs_input_grad)
        xin = xs[0].detach() if n_in == 1 else tuple(x.detach() for x in xs)
        out, tape = program.forward(xin, P, buffers, training, need_any, extra)
        # The tape may hold the output tensor itself (e.g. the generator's tanh output, needed by its backward).  The
        # object returned from here gets grad_fn = this node, so returning the SAME object would close a reference
        # cycle node -> tape -> output -> node and keep the whole tape (1.7 GiB for the generator at batch 8) alive until
        # the cyclic garbage collector runs.  Return an alias instead.
        out = tuple(o.detach() for o in out) if isinstance(out, (tuple, list)) else out.detach()
        ctx.program = program
        ctx.tape = tape if need_any else None
        ctx.names = names
        ctx.n_in = n_in
        ctx.param_refs = params if need_any else None
        return out

    @staticmethod
    def backward(ctx, dout):
        if ctx.tape is None:
            raise RuntimeError("genseg_b200: backward called but no tape was saved")
        n_in = ctx.n_in
        need_in = ctx.needs_input_grad  # (program, buffers, training, extra, n_in, grad_mode, *inputs, *params)
        need = {k: bool(need_in[6 + n_in + i]) for i, k in enumerate(ctx.names)}
        need_x = [bool(need_in[6 + i]) for i in range(n_in)]
        dx, G = ctx.program.backward(ctx.tape, dout, need, need_x[0] if n_in == 1 else need_x)
        dxs = (dx,) if n_in == 1 else tuple(dx)
        dxs = tuple(d if nx else None for d, nx in zip(dxs, need_x))
        grads = [G.get(k) if need[k] else None for k in ctx.names]
        if G.get("__reduced__") and ctx.param_refs is not None:
            # data parallelism: the program summed these gradients over the ranks inside backward (parallel.py)
            for p_, g_ in zip(ctx.param_refs, grads):
                if g_ is not None:
                    p_._gsb_grad_reduced = True
        return (None, None, None, None, None, None, *dxs, *grads)


def run_program(program, module: torch.nn.Module, x, extra=None):
    """Collect the module's parameters/buffers by state_dict name and run the program."""
    xs = x if isinstance(x, (tuple, list)) else (x,)
    for t in xs:
        if not t.is_cuda:
            raise RuntimeError("genseg_b200: the B200 hot path needs CUDA tensors; there is no CPU fallback "
                               "(got a %s tensor)" % t.device.type)
    pd = dict(module.named_parameters())
    bd = dict(module.named_buffers())
    params = [pd[k] for k in program.param_names]
    return ProgramFunction.apply(program, bd, module.training, extra, len(xs), torch.is_grad_enabled(), *xs, *params)
