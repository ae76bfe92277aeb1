"""CPU oracle for the GenSeg hot path  --  TEST INFRASTRUCTURE ONLY.

This file is a *restatement* (not a copy) of the arithmetic that the reference
(importZL/semantic_segmentation, mounted read-only at /root/reference while the
repo is being built) performs on its U-Net / Pix2Pix hot path, written as plain
functional fp32 PyTorch-on-CPU plus numpy for the integer Dice counts.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s CPU-baseline leg
may import it.  The product package (``semantic_segmentation_b200``) never does:
its CUDA path fails loudly when the extension is missing.

Parity pinning: the reference ships no tests and no golden vectors (SURVEY.md
section 4/8c).  The oracle is therefore pinned against *outputs of the reference
itself*: ``tests/golden/make_golden.py`` imports the unmodified reference
modules from /root/reference, runs them on seeded inputs/weights produced by
``seeded_state`` / ``seeded_inputs`` below, and commits the results as
``tests/golden/*.npz``.  ``tests/test_oracle_golden.py`` checks this file
against those fixtures on CPU.

The third-party arithmetic on the path is PyTorch's (reference pins
torch==1.13.1, env.sh:3; this image has torch 2.11): conv2d, conv_transpose2d,
batch_norm, max_pool2d, bilinear interpolate, cross_entropy,
binary_cross_entropy_with_logits.  The oracle calls those same CPU functionals.

Reference citations (file:line, relative to the reference root) are given on
every function.
"""
from __future__ import annotations

import math
from collections import OrderedDict
from typing import Dict, List, Optional, Tuple

import numpy as np
import torch
import torch.nn.functional as F

Tensor = torch.Tensor
BN_EPS = 1e-5
BN_MOMENTUM = 0.1


# ----------------------------------------------------------------------------
# Seeded weights / inputs shared by the golden generator and the tests
# ----------------------------------------------------------------------------

def unet_state_spec(n_channels: int, n_classes: int, bilinear: bool) -> "OrderedDict[str, Tuple[int, ...]]":
    """Key -> shape of ``unet.UNet(n_channels, n_classes, bilinear).state_dict()``.

    Mirrors unet/unet_model.py:8-24 and unet/unet_parts.py:8-77.
    """
    spec: "OrderedDict[str, Tuple[int, ...]]" = OrderedDict()

    def bn(prefix: str, c: int) -> None:
        spec[prefix + ".weight"] = (c,)
        spec[prefix + ".bias"] = (c,)
        spec[prefix + ".running_mean"] = (c,)
        spec[prefix + ".running_var"] = (c,)
        spec[prefix + ".num_batches_tracked"] = ()

    def dconv(prefix: str, cin: int, cout: int, mid: Optional[int] = None) -> None:
        mid = mid or cout
        spec[prefix + ".double_conv.0.weight"] = (mid, cin, 3, 3)
        bn(prefix + ".double_conv.1", mid)
        spec[prefix + ".double_conv.3.weight"] = (cout, mid, 3, 3)
        bn(prefix + ".double_conv.4", cout)

    factor = 2 if bilinear else 1
    dconv("inc", n_channels, 64)
    dconv("down1.maxpool_conv.1", 64, 128)
    dconv("down2.maxpool_conv.1", 128, 256)
    dconv("down3.maxpool_conv.1", 256, 512)
    dconv("down4.maxpool_conv.1", 512, 1024 // factor)
    for name, cin, cout in (("up1", 1024, 512 // factor), ("up2", 512, 256 // factor),
                            ("up3", 256, 128 // factor), ("up4", 128, 64)):
        if bilinear:
            dconv(name + ".conv", cin, cout, cin // 2)
        else:
            spec[name + ".up.weight"] = (cin, cin // 2, 2, 2)
            spec[name + ".up.bias"] = (cin // 2,)
            dconv(name + ".conv", cin, cout)
    spec["outc.conv.weight"] = (n_classes, 64, 1, 1)
    spec["outc.conv.bias"] = (n_classes,)
    return spec


def generator_state_spec(input_nc: int, output_nc: int, ngf: int = 64, num_downs: int = 8
                         ) -> "OrderedDict[str, Tuple[int, ...]]":
    """Key -> shape of the ``unet_256`` (num_downs 8) / ``unet_128`` (num_downs 7, networks.py:148-151) generator built
    by models_pix2pix/networks.py:514-617 with BatchNorm (use_bias False except the
    outermost up-cell, :590).  ``layer_index`` 0 = innermost."""
    spec: "OrderedDict[str, Tuple[int, ...]]" = OrderedDict()

    def bn(prefix: str, c: int) -> None:
        spec[prefix + ".weight"] = (c,)
        spec[prefix + ".bias"] = (c,)
        spec[prefix + ".running_mean"] = (c,)
        spec[prefix + ".running_var"] = (c,)
        spec[prefix + ".num_batches_tracked"] = ()

    # (outer_nc, inner_nc) from the outermost block inwards (networks.py:532-546)
    levels = [(output_nc, ngf), (ngf, ngf * 2), (ngf * 2, ngf * 4), (ngf * 4, ngf * 8)] + [(ngf * 8, ngf * 8)] * (num_downs - 4)
    prefix = "model.model"
    # outermost: [downconv, submodule, uprelu, upcell, tanh]
    spec[prefix + ".0.weight"] = (ngf, input_nc, 4, 4)
    tail: List[Tuple[str, Tuple[int, ...]]] = []
    for k in (4, 6, 8):
        i = (k - 4) // 2
        tail.append((f"{prefix}.3._ops._ops.{i}.op.weight", (ngf * 2, output_nc, k, k)))
        tail.append((f"{prefix}.3._ops._ops.{i}.op.bias", (output_nc,)))
    tails = [tail]
    p = prefix + ".1"
    for depth in range(1, num_downs):
        outer, inner = levels[depth]
        innermost = depth == num_downs - 1
        spec[p + ".model.1.weight"] = (inner, outer, 4, 4)
        t: List[Tuple[str, Tuple[int, ...]]] = []
        if innermost:
            # [downrelu, downconv, uprelu, upcell, upnorm]
            for k in (4, 6, 8):
                t.append((f"{p}.model.3._ops._ops.{(k - 4) // 2}.op.weight", (inner, outer, k, k)))
            for s in ("weight", "bias", "running_mean", "running_var"):
                t.append((f"{p}.model.4.{s}", (outer,)))
            t.append((f"{p}.model.4.num_batches_tracked", ()))
        else:
            # [downrelu, downconv, downnorm, submodule, uprelu, upcell, upnorm, (dropout)]
            bn(p + ".model.2", inner)
            for k in (4, 6, 8):
                t.append((f"{p}.model.5._ops._ops.{(k - 4) // 2}.op.weight", (inner * 2, outer, k, k)))
            for s in ("weight", "bias", "running_mean", "running_var"):
                t.append((f"{p}.model.6.{s}", (outer,)))
            t.append((f"{p}.model.6.num_batches_tracked", ()))
        tails.append(t)
        p = p + ".model.3"
    for t in reversed(tails):
        for k_, s_ in t:
            spec[k_] = s_
    return spec


def discriminator_state_spec(input_nc: int, ndf: int = 64, n_layers: int = 3) -> "OrderedDict[str, Tuple[int, ...]]":
    """Key -> shape of the PatchGAN, networks.py:620-665 ('basic' = 3 layers; 'n_layers' takes --n_layers_D)."""
    spec: "OrderedDict[str, Tuple[int, ...]]" = OrderedDict()
    spec["model.0.weight"] = (ndf, input_nc, 4, 4)
    spec["model.0.bias"] = (ndf,)
    chans = [ndf * min(2 ** i, 8) for i in range(n_layers + 1)]
    idx = 2
    for i in range(1, n_layers + 1):
        spec[f"model.{idx}.weight"] = (chans[i], chans[i - 1], 4, 4)
        for s in ("weight", "bias", "running_mean", "running_var"):
            spec[f"model.{idx + 1}.{s}"] = (chans[i],)
        spec[f"model.{idx + 1}.num_batches_tracked"] = ()
        idx += 3
    spec[f"model.{idx}.weight"] = (1, chans[n_layers], 4, 4)
    spec[f"model.{idx}.bias"] = (1,)
    return spec


def seeded_state(spec: "OrderedDict[str, Tuple[int, ...]]", seed: int, conv_std: Optional[float] = None
                 ) -> "OrderedDict[str, Tensor]":
    """Deterministic fp32 state_dict for a key->shape spec (test fixture weights).

    conv / convT weights: N(0, conv_std) or, if conv_std is None, N(0, sqrt(2/fan_in))
    so activations keep O(1) scale through the U-Net; BN weight ~ 1 + 0.1 N(0,1),
    BN bias ~ 0.1 N(0,1); conv bias ~ 0.1 N(0,1); running_mean 0.05 N, running_var
    in [0.5, 1.5]; num_batches_tracked 0.  Values depend only on (key order,
    shape, seed): one CPU torch.Generator stream consumed in key order.
    """
    g = torch.Generator().manual_seed(seed)
    out: "OrderedDict[str, Tensor]" = OrderedDict()
    for k, shp in spec.items():
        if k.endswith("num_batches_tracked"):
            out[k] = torch.zeros((), dtype=torch.int64)
        elif k.endswith("running_mean"):
            out[k] = 0.05 * torch.randn(shp, generator=g)
        elif k.endswith("running_var"):
            out[k] = 0.5 + torch.rand(shp, generator=g)
        elif len(shp) == 4:
            fan_in = shp[1] * shp[2] * shp[3]
            if "_ops" in k or ".up.weight" in k:  # ConvTranspose2d: [Cin, Cout, k, k], stride 2
                fan_in = shp[0] * shp[2] * shp[3] / 4.0
            std = conv_std if conv_std is not None else math.sqrt(2.0 / fan_in)
            out[k] = std * torch.randn(shp, generator=g)
        elif k.endswith(".bias") and (".op." in k or "conv." in k or ".up." in k
                                      or k in ("model.0.bias", "model.11.bias")):
            out[k] = 0.1 * torch.randn(shp, generator=g)
        elif k.endswith(".weight"):
            out[k] = 1.0 + 0.1 * torch.randn(shp, generator=g)
        else:
            out[k] = 0.1 * torch.randn(shp, generator=g)
    return out


def seeded_inputs(seed: int, n: int, c: int, h: int, w: int, n_classes: int) -> Tuple[Tensor, Tensor]:
    """Synthetic JSRT-shaped batch: image in [0,1) fp32 [n,c,h,w]; int64 labels
    ([n,h,w] for n_classes>1, [n,1,h,w] for n_classes==1, cf.
    running_files/train_end2end_jsrt.py:182)."""
    g = torch.Generator().manual_seed(seed)
    x = torch.rand(n, c, h, w, generator=g)
    hi = max(n_classes, 2)
    if n_classes == 1:
        y = torch.randint(0, hi, (n, 1, h, w), generator=g)
    else:
        y = torch.randint(0, hi, (n, h, w), generator=g)
    return x, y


# ----------------------------------------------------------------------------
# U-Net (unet/unet_parts.py, unet/unet_model.py)
# ----------------------------------------------------------------------------

def _bn(x: Tensor, st: Dict[str, Tensor], prefix: str, training: bool, update: bool) -> Tensor:
    """nn.BatchNorm2d(momentum 0.1, eps 1e-5): unet/unet_parts.py:17,20 and
    networks.py:32-33.  Train: batch mean / biased var; running stats updated with
    the unbiased var (in place in ``st`` when ``update``)."""
    rm, rv = st[prefix + ".running_mean"], st[prefix + ".running_var"]
    if training and not update:
        rm, rv = rm.clone(), rv.clone()
    y = F.batch_norm(x, rm, rv, st[prefix + ".weight"], st[prefix + ".bias"],
                     training=training, momentum=BN_MOMENTUM, eps=BN_EPS)
    if training and update:
        st[prefix + ".num_batches_tracked"] += 1
    return y


def double_conv(x: Tensor, st: Dict[str, Tensor], prefix: str, training: bool, update: bool,
                taps: Optional[Dict[str, Tensor]] = None) -> Tensor:
    """DoubleConv, unet/unet_parts.py:8-25: (Conv3x3 p1 no-bias -> BN -> ReLU) x 2."""
    p = prefix + ".double_conv"
    x = F.conv2d(x, st[p + ".0.weight"], None, padding=1)
    if taps is not None:
        taps[p + ".0"] = x
    x = F.relu(_bn(x, st, p + ".1", training, update))
    x = F.conv2d(x, st[p + ".3.weight"], None, padding=1)
    if taps is not None:
        taps[p + ".3"] = x
    x = F.relu(_bn(x, st, p + ".4", training, update))
    return x


def up_block(x1: Tensor, x2: Tensor, st: Dict[str, Tensor], name: str, bilinear: bool, training: bool,
             update: bool, taps: Optional[Dict[str, Tensor]] = None) -> Tensor:
    """Up, unet/unet_parts.py:42-68: upsample x1 (bilinear align_corners=True, or
    ConvTranspose2d k2 s2), zero-pad to x2's size (left = diff//2), cat([x2, x1])."""
    if bilinear:
        x1 = F.interpolate(x1, scale_factor=2, mode="bilinear", align_corners=True)
    else:
        x1 = F.conv_transpose2d(x1, st[name + ".up.weight"], st[name + ".up.bias"], stride=2)
    dy = x2.shape[2] - x1.shape[2]
    dx = x2.shape[3] - x1.shape[3]
    x1 = F.pad(x1, [dx // 2, dx - dx // 2, dy // 2, dy - dy // 2])
    x = torch.cat([x2, x1], dim=1)
    return double_conv(x, st, name + ".conv", training, update, taps)


def unet_forward(st: Dict[str, Tensor], x: Tensor, bilinear: bool, training: bool = True,
                 update_running: bool = False, taps: Optional[Dict[str, Tensor]] = None) -> Tensor:
    """UNet.forward, unet/unet_model.py:26-37.  ``taps`` collects the raw conv
    outputs (pre-BN) and the block outputs for per-layer parity checks."""
    x1 = double_conv(x, st, "inc", training, update_running, taps)
    xs = [x1]
    for i in range(1, 5):
        xs.append(double_conv(F.max_pool2d(xs[-1], 2), st, f"down{i}.maxpool_conv.1", training,
                              update_running, taps))
    h = xs[4]
    for i, skip in zip(range(1, 5), (xs[3], xs[2], xs[1], xs[0])):
        h = up_block(h, skip, st, f"up{i}", bilinear, training, update_running, taps)
        if taps is not None:
            taps[f"up{i}"] = h
    if taps is not None:
        for i, t in enumerate(xs):
            taps[f"x{i + 1}"] = t
    return F.conv2d(h, st["outc.conv.weight"], st["outc.conv.bias"])  # OutConv, unet_parts.py:71-77


# ----------------------------------------------------------------------------
# Dice / losses (util/dice_score.py, running_files/train_end2end_jsrt.py:136,182-183)
# ----------------------------------------------------------------------------

def dice_coeff(inp: Tensor, tgt: Tensor, reduce_batch_first: bool = False, epsilon: float = 1e-6) -> Tensor:
    """util/dice_score.py:5-17."""
    assert inp.size() == tgt.size()
    assert inp.dim() == 3 or not reduce_batch_first
    dims = (-1, -2) if inp.dim() == 2 or not reduce_batch_first else (-1, -2, -3)
    inter = 2 * (inp * tgt).sum(dim=dims)
    sets = inp.sum(dim=dims) + tgt.sum(dim=dims)
    sets = torch.where(sets == 0, inter, sets)
    return ((inter + epsilon) / (sets + epsilon)).mean()


def multiclass_dice_coeff(inp: Tensor, tgt: Tensor, reduce_batch_first: bool = False,
                          epsilon: float = 1e-6) -> Tensor:
    """util/dice_score.py:20-22."""
    return dice_coeff(inp.flatten(0, 1), tgt.flatten(0, 1), reduce_batch_first, epsilon)


def dice_loss(inp: Tensor, tgt: Tensor, multiclass: bool = False) -> Tensor:
    """util/dice_score.py:25-28 (global-batch ratio: reduce_batch_first=True)."""
    fn = multiclass_dice_coeff if multiclass else dice_coeff
    return 1 - fn(inp, tgt, reduce_batch_first=True)


def seg_loss(logits: Tensor, target: Tensor) -> Tuple[Tensor, Tensor, Tensor]:
    """Training loss of the segmentation step -> (total, ce_or_bce, dice).

    n_classes == 1: BCEWithLogits(pred, y.float()) + dice_loss(sigmoid(pred.squeeze(1)),
    y.float().squeeze(1))  (running_files/train_end2end_jsrt.py:136,182-183).
    n_classes > 1: CrossEntropy(pred, y) + dice_loss(softmax(pred,1), one_hot(y),
    multiclass=True) (composition of :136 with util/dice_score.py:25-28).
    """
    c = logits.shape[1]
    if c == 1:
        t = target.float()
        ce = F.binary_cross_entropy_with_logits(logits, t)
        d = dice_loss(torch.sigmoid(logits.squeeze(1)), t.squeeze(1), multiclass=False)
    else:
        ce = F.cross_entropy(logits, target)
        oh = F.one_hot(target, c).permute(0, 3, 1, 2).float()
        d = dice_loss(F.softmax(logits, dim=1).float(), oh, multiclass=True)
    return ce + d, ce, d


def eval_counts(logits: np.ndarray, target: np.ndarray) -> Dict[str, np.ndarray]:
    """Integer Dice counts of unet/evaluate.py:29-40 from given logits (numpy).

    n_classes == 1: pred = sigmoid(x) > 0.5 computed in fp32 exactly like torch
    (1/(1+exp(-x)) rounded to fp32, compared with 0.5); counts per image.
    n_classes > 1: pred = argmax over channels (first max wins; NaN counts as max),
    counts per (image, class) for classes 1..C-1 (background dropped, :40).
    Returns int64 arrays inter = sum(pred*true), psum = sum(pred), tsum = sum(true),
    plus the uint8 mask.
    """
    n, c = logits.shape[0], logits.shape[1]
    if c == 1:
        x = logits[:, 0].astype(np.float32)
        sig = torch.sigmoid(torch.from_numpy(x)).numpy()
        pred = (sig > np.float32(0.5))
        true = target.reshape(n, *x.shape[1:]).astype(np.int64)
        inter = (pred & (true == 1)).reshape(n, -1).sum(1).astype(np.int64)[:, None]
        psum = pred.reshape(n, -1).sum(1).astype(np.int64)[:, None]
        tsum = (true == 1).reshape(n, -1).sum(1).astype(np.int64)[:, None]
        return dict(inter=inter, psum=psum, tsum=tsum, mask=pred.astype(np.uint8))
    am = torch.from_numpy(logits.astype(np.float32)).argmax(dim=1).numpy()
    true = target.reshape(n, *am.shape[1:]).astype(np.int64)
    inter = np.zeros((n, c - 1), np.int64)
    psum = np.zeros((n, c - 1), np.int64)
    tsum = np.zeros((n, c - 1), np.int64)
    for k in range(1, c):
        p, t = am == k, true == k
        inter[:, k - 1] = (p & t).reshape(n, -1).sum(1)
        psum[:, k - 1] = p.reshape(n, -1).sum(1)
        tsum[:, k - 1] = t.reshape(n, -1).sum(1)
    return dict(inter=inter, psum=psum, tsum=tsum, mask=am.astype(np.uint8))


def dice_from_counts(inter: np.ndarray, psum: np.ndarray, tsum: np.ndarray, epsilon: float = 1e-6) -> float:
    """dice_coeff(reduce_batch_first=False) evaluated on integer counts in fp32,
    util/dice_score.py:12-17 (empty-set rule: sets==0 -> inter)."""
    i2 = (2 * inter).astype(np.float32)
    s = (psum + tsum).astype(np.float32)
    s = np.where(s == 0, i2, s)
    return float(((i2 + np.float32(epsilon)) / (s + np.float32(epsilon))).astype(np.float32).mean(dtype=np.float32))


# ----------------------------------------------------------------------------
# Pix2Pix generator / discriminator / losses (models_pix2pix/networks.py)
# ----------------------------------------------------------------------------

def mixed_upconv(x: Tensor, st: Dict[str, Tensor], prefix: str, alpha_row: Tensor, bias: bool) -> Tensor:
    """Cell_upconv / MixedOp_upconv, networks.py:486-511 with the primitives of
    architecture_pix2pix/operations.py:14-39: sum_i softmax(alpha)_i * ConvT_{k_i}(x),
    k in (4,6,8), stride 2, padding (k-2)/2.  Executed un-merged, as the reference does."""
    w = F.softmax(alpha_row, dim=-1)
    out = 0
    for i, k in enumerate((4, 6, 8)):
        b = st[f"{prefix}._ops._ops.{i}.op.bias"] if bias else None
        out = out + w[i] * F.conv_transpose2d(x, st[f"{prefix}._ops._ops.{i}.op.weight"], b,
                                              stride=2, padding=(k - 2) // 2)
    return out


def merged_upconv_weight(st: Dict[str, Tensor], prefix: str, alpha_row: Tensor, bias: bool
                         ) -> Tuple[Tensor, Optional[Tensor]]:
    """The single 8x8/s2/p3 ConvTranspose2d kernel equivalent to ``mixed_upconv``
    (SURVEY fact 4): W = w2*W8 + pad1(w1*W6) + pad2(w0*W4); bias = sum_i w_i b_i."""
    w = F.softmax(alpha_row, dim=-1)
    W = w[2] * st[f"{prefix}._ops._ops.2.op.weight"]
    W = W + F.pad(w[1] * st[f"{prefix}._ops._ops.1.op.weight"], [1, 1, 1, 1])
    W = W + F.pad(w[0] * st[f"{prefix}._ops._ops.0.op.weight"], [2, 2, 2, 2])
    b = None
    if bias:
        b = sum(w[i] * st[f"{prefix}._ops._ops.{i}.op.bias"] for i in range(3))
    return W, b


def generator_forward(st: Dict[str, Tensor], x: Tensor, upconv_arch: Tensor, training: bool = True,
                      update_running: bool = False, dropout_masks: Optional[List[Tensor]] = None,
                      taps: Optional[Dict[str, Tensor]] = None, num_downs: int = 8) -> Tensor:
    """UnetGenerator('unet_256' | 'unet_128').forward, networks.py:514-617 (num_downs 8 | 7).

    Restated iteratively.  Level d = 0 is the outermost block (layer_index num_downs-1), d = num_downs-1
    the innermost (layer_index 0).  The text below is for num_downs = 8; with 7 the dropout blocks are depth 4,5.  The in-place LeakyReLU (networks.py:583) makes
    each skip carry lrelu(x).  Dropout(0.5) follows the up-norm of the
    blocks at depth 4,5,6 (layer_index 3,2,1; networks.py:535-537,606-607);
    ``dropout_masks`` = optional list of three {0,1} keep-masks ordered by depth
    (4,5,6) applied as x*mask*2; None with training=True means dropout disabled
    (parity mode).
    """
    p = "model.model"
    e = F.conv2d(x, st[p + ".0.weight"], None, stride=2, padding=1)  # outermost downconv, no norm
    skips: List[Tensor] = []
    prefixes: List[str] = []
    q = p + ".1"
    last = num_downs - 1
    for d in range(1, num_downs):
        s = F.leaky_relu(e, 0.2)
        skips.append(s)
        prefixes.append(q)
        e = F.conv2d(s, st[q + ".model.1.weight"], None, stride=2, padding=1)
        if d != last:
            e = _bn(e, st, q + ".model.2", training, update_running)
        if taps is not None:
            taps[f"down{d}"] = e
        q = q + ".model.3"
    # innermost up: relu -> mixed convT -> BN
    h = F.relu(e)
    for d in range(last, 0, -1):
        q = prefixes[d - 1]
        up_idx, bn_idx = (3, 4) if d == last else (5, 6)
        layer_index = last - d
        h = mixed_upconv(h, st, f"{q}.model.{up_idx}", upconv_arch[layer_index], bias=False)
        h = _bn(h, st, f"{q}.model.{bn_idx}", training, update_running)
        if 4 <= d < last and dropout_masks is not None:
            h = h * dropout_masks[d - 4] * 2.0
        if taps is not None:
            taps[f"up{d}"] = h
        h = F.relu(torch.cat([skips[d - 1], h], dim=1))  # parent's uprelu over the concat
    out = mixed_upconv(h, st, p + ".3", upconv_arch[last], bias=True)
    return torch.tanh(out)


def discriminator_forward(st: Dict[str, Tensor], x: Tensor, training: bool = True,
                          update_running: bool = False, taps: Optional[Dict[str, Tensor]] = None) -> Tensor:
    """NLayerDiscriminator.forward, networks.py:620-665 (n_layers read off the state: 3 for 'basic')."""
    n_layers = sum(1 for k in st if k.endswith(".running_mean"))
    h = F.leaky_relu(F.conv2d(x, st["model.0.weight"], st["model.0.bias"], stride=2, padding=1), 0.2)
    for idx, stride in [(2 + 3 * i, 2 if i < n_layers - 1 else 1) for i in range(n_layers)]:
        h = F.conv2d(h, st[f"model.{idx}.weight"], None, stride=stride, padding=1)
        if taps is not None:
            taps[f"conv{idx}"] = h
        h = F.leaky_relu(_bn(h, st, f"model.{idx + 1}", training, update_running), 0.2)
    idx = 2 + 3 * n_layers
    return F.conv2d(h, st[f"model.{idx}.weight"], st[f"model.{idx}.bias"], stride=1, padding=1)


def gan_loss(pred: Tensor, target_is_real: bool, gan_mode: str = "vanilla") -> Tensor:
    """GANLoss.__call__, networks.py:263-281 (labels exactly 1.0 / 0.0)."""
    if gan_mode == "wgangp":
        return -pred.mean() if target_is_real else pred.mean()
    t = torch.full_like(pred, 1.0 if target_is_real else 0.0)
    if gan_mode == "vanilla":
        return F.binary_cross_entropy_with_logits(pred, t)
    if gan_mode == "lsgan":
        return F.mse_loss(pred, t)
    raise NotImplementedError(gan_mode)


def pix2pix_losses(stG: Dict[str, Tensor], stD: Dict[str, Tensor], arch: Tensor, mask: Tensor, image: Tensor,
                   lambda_l1: float = 100.0, training_g: bool = True) -> Dict[str, Tensor]:
    """The arithmetic of Pix2PixModel.optimize_parameters without the optimiser
    (models_pix2pix/pix2pix_model.py:102-130): returns fake, loss_D, loss_G parts."""
    fake = generator_forward(stG, mask, arch, training=training_g)
    pred_fake_d = discriminator_forward(stD, torch.cat([mask, fake.detach()], 1))
    pred_real = discriminator_forward(stD, torch.cat([mask, image], 1))
    loss_d = 0.5 * (gan_loss(pred_fake_d, False) + gan_loss(pred_real, True))
    pred_fake = discriminator_forward(stD, torch.cat([mask, fake], 1))
    g_gan = gan_loss(pred_fake, True)
    g_l1 = F.l1_loss(fake, image) * lambda_l1
    return dict(fake=fake, pred_fake=pred_fake, pred_real=pred_real, loss_D=loss_d, G_GAN=g_gan,
                G_L1=g_l1, loss_G=g_gan + g_l1)
