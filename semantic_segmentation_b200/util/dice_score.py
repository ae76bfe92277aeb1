"""Dice coefficient / loss -- drop-in for the reference's util/dice_score.py:5-28 (same names, signatures,
assertions and epsilon / empty-set rule), with the reductions on the sm_100a kernels.

``seg_loss`` is the fused training loss of the hot path: CrossEntropy (or BCE-with-logits) + Dice loss in
two bandwidth passes over the logits (running_files/train_end2end_jsrt.py:136,182-183).
"""
import torch
from torch import Tensor

from .. import engine, ops


class _DiceCoeff(torch.autograd.Function):
    @staticmethod
    def forward(ctx, inp: Tensor, tgt: Tensor, slices: int, inner: int, epsilon: float):
        a = inp.contiguous().float()
        b = tgt.contiguous().float()
        sums = ops.dice_sums(a, b, slices, inner)            # [slices, 3] fp64: sum a*b, sum a, sum b
        inter = (2.0 * sums[:, 0]).float()
        raw = (sums[:, 1] + sums[:, 2]).float()
        empty = raw == 0                                     # util/dice_score.py:14: sets_sum[sets_sum == 0] = inter
        sets = torch.where(empty, inter, raw)
        dice = (inter + epsilon) / (sets + epsilon)
        ctx.save_for_backward(a, b, inter, sets, empty)
        ctx.meta = (slices, inner, epsilon, inp.shape, inp.dtype, tgt.shape, tgt.dtype)
        return dice.mean()                                   # no host synchronisation: graph-capturable

    @staticmethod
    def backward(ctx, g):
        a, b, inter, sets, empty = ctx.saved_tensors
        slices, inner, eps, shape, dtype, tshape, tdtype = ctx.meta
        # dice_s = (2I+eps)/(S+eps), I = sum a*b, S = sum a + sum b:
        #   d dice_s / d a = (2 b (S+eps) - (2I+eps)) / (S+eps)^2, and symmetrically for b.
        # An empty slice (S == 0) has dice_s = (2I+eps)/(2I+eps) == 1 identically: zero gradient, like the reference's graph.
        den = (sets + eps)
        keep = (~empty).to(den.dtype)
        ca = (2.0 / den) * (g / slices) * keep
        cb = ((inter + eps) / (den * den)) * (g / slices) * keep
        da = db = None
        if ctx.needs_input_grad[0]:
            da = (b.view(slices, inner) * ca[:, None] - cb[:, None]).view(shape).to(dtype)
        if ctx.needs_input_grad[1] and tdtype.is_floating_point:
            db = (a.view(slices, inner) * ca[:, None] - cb[:, None]).view(tshape).to(tdtype)
        return da, db, None, None, None


def dice_coeff(input: Tensor, target: Tensor, reduce_batch_first: bool = False, epsilon: float = 1e-6):
    # Average of Dice coefficient for all batches, or for a single mask
    assert input.size() == target.size()
    assert input.dim() == 3 or not reduce_batch_first
    if not input.is_cuda:
        raise RuntimeError("genseg_b200: dice_coeff needs CUDA tensors; there is no CPU fallback")
    if input.dim() == 2 or not reduce_batch_first:
        inner = input.shape[-1] * input.shape[-2]
        slices = input.numel() // inner
    else:
        inner = input.shape[-1] * input.shape[-2] * input.shape[-3]
        slices = input.numel() // inner
    return _DiceCoeff.apply(input, target, slices, inner, epsilon)


def multiclass_dice_coeff(input: Tensor, target: Tensor, reduce_batch_first: bool = False, epsilon: float = 1e-6):
    # Average of Dice coefficient for all classes
    return dice_coeff(input.flatten(0, 1), target.flatten(0, 1), reduce_batch_first, epsilon)


def dice_loss(input: Tensor, target: Tensor, multiclass: bool = False):
    # Dice loss (objective to minimize) between 0 and 1
    fn = multiclass_dice_coeff if multiclass else dice_coeff
    return 1 - fn(input, target, reduce_batch_first=True)


class _SegLoss(torch.autograd.Function):
    @staticmethod
    def forward(ctx, logits: Tensor, labels: Tensor):
        lg = logits.contiguous()
        if lg.dtype != torch.float32:
            lg = lg.float()
        lb = labels.contiguous()
        if lb.dtype != torch.int64:
            lb = lb.long()
        n, c, h, w = lg.shape
        if lb.numel() != n * h * w:
            raise RuntimeError("seg_loss: labels must hold one class index per pixel")
        sums = ops.seg_loss_reduce(lg, lb)
        count = float(n * h * w)
        red = engine._REDUCER
        if red["fn"] is not None and red["world"] > 1:   # global-batch Dice / CE under data parallelism
            red["fn"](sums)
            count *= red["world"]
        out = ops.seg_loss_finalize(sums, count)
        ctx.save_for_backward(lg, lb, sums)
        ctx.count = count
        total, ce, dl = out[0].clone(), out[1].clone(), out[2].clone()
        ctx.mark_non_differentiable(ce, dl)
        return total, ce, dl

    @staticmethod
    def backward(ctx, g, _g1, _g2):
        lg, lb, sums = ctx.saved_tensors
        gs = g.contiguous().float()
        return ops.seg_loss_bwd(lg, lb, sums, ctx.count, gs), None


def seg_loss(logits: Tensor, labels: Tensor, return_parts: bool = False):
    """Fused CE/BCE + Dice training loss (n_classes from logits.shape[1]).

    n_classes == 1: BCEWithLogits(pred, y.float()) + dice_loss(sigmoid(pred.squeeze(1)), y.float().squeeze(1));
    n_classes  > 1: CrossEntropy(pred, y) + dice_loss(softmax(pred, 1), one_hot(y), multiclass=True).
    Under parallel.init_data_parallel() the Dice ratio and CE mean are over the GLOBAL batch."""
    if not logits.is_cuda:
        raise RuntimeError("genseg_b200: seg_loss needs CUDA tensors; there is no CPU fallback")
    total, ce, dl = _SegLoss.apply(logits, labels)
    return (total, ce, dl) if return_parts else total
