"""Layer executor of the B200 hot path: explicit forward / backward over NHWC device buffers.

Autograd sees ONE node per network call (see ``functions.py``); inside it every kernel launch,
saved activation and gradient buffer is managed here, so skip gradients accumulate in place,
concat buffers are written by their producers, and nothing runs through ATen / cuDNN.
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Dict, List, Optional, Tuple

import torch

from . import ops
from ._lib import (ACT_LRELU, ACT_NONE, ACT_RELU, ACT_TANH, MODE_CONV, MODE_CONVT, GsbView, view_nchw, view_nhwc)

BN_EPS = 1e-5
BN_MOMENTUM = 0.1

_CONFIG = {"dtype": torch.bfloat16, "use_tc": 2, "fused_stats": True}


def set_precision(mode: str) -> None:
    """'bf16' (production: bf16 activations/weights, fp32 accumulate, tcgen05 kernels) or
    'fp32' (validation mode: fp32 activations/weights through the SIMT kernels)."""
    if mode == "bf16":
        _CONFIG["dtype"] = torch.bfloat16
    elif mode == "fp32":
        _CONFIG["dtype"] = torch.float32
    else:
        raise ValueError("precision must be 'bf16' or 'fp32'")


def get_precision() -> str:
    return "bf16" if _CONFIG["dtype"] == torch.bfloat16 else "fp32"


def set_tensor_core_mode(use_tc: int, fused_stats: bool = True) -> None:
    """0 = SIMT kernels only, 1 = require tcgen05 where requested, 2 = auto (default)."""
    _CONFIG["use_tc"] = int(use_tc)
    _CONFIG["fused_stats"] = bool(fused_stats)


def act_dtype() -> torch.dtype:
    return _CONFIG["dtype"]


# --- cross-rank reductions (sync-BN / global Dice); installed by parallel.py ----------------------
_REDUCER = {"fn": None, "world": 1, "rank": 0, "sync_bn": False, "peer": None, "grad_fn": None, "sm_reserve": 0}


def first_global_sample(n_local: int) -> int:
    """Index of this rank's first sample in the global batch (equal shards, rank-major: parallel.shard_batch)."""
    return int(_REDUCER.get("rank", 0)) * int(n_local) if int(_REDUCER.get("world", 1)) > 1 else 0


def centered_bn() -> bool:
    """fp32 validation mode: BatchNorm is applied as (x - mean) * scale + beta like torch; the bf16 production mode keeps the
    folded x * scale + shift (its inputs carry 2^-9 relative rounding anyway)."""
    return _CONFIG["dtype"] == torch.float32


def install_reducer(fn, world: int, sync_bn: bool, peer=None, grad_fn=None, sm_reserve: int = 0, rank: int = 0) -> None:
    """peer = {"regions", "rank", "world"}: NVLink peer regions (parallel.py) -- the BatchNorm statistic kernels then do
    their cross-rank sums themselves instead of calling ``fn`` between launches.
    grad_fn(tensor) -> handle with .wait(): asynchronous in-place gradient all-reduce; when set, the U-Net program
    all-reduces its gradients inside backward (first bucket overlapped with the encoder backward)."""
    _REDUCER.update(fn=fn, world=int(world), sync_bn=bool(sync_bn), peer=peer, grad_fn=grad_fn, sm_reserve=int(sm_reserve),
                    rank=int(rank))


def _fused_stats_mode():
    """(usable, peer, world): the fused statistic kernels cover one GPU and sync-BN over peer memory; sync-BN through a
    plain reducer (gloo / NCCL fallback) keeps the separate reduce -> all-reduce -> finalize launches."""
    if not _CONFIG.get("fused_bn_stats", True):
        return False, None, 1
    if _REDUCER["sync_bn"] and _REDUCER["fn"] is not None and _REDUCER["world"] > 1:
        if _REDUCER["peer"] is None:
            return False, None, 1
        return True, _REDUCER["peer"], _REDUCER["world"]
    return True, None, 1


def _sync_stats(sums: torch.Tensor, count: float) -> Tuple[torch.Tensor, float]:
    if _REDUCER["sync_bn"] and _REDUCER["fn"] is not None and _REDUCER["world"] > 1:
        _REDUCER["fn"](sums)
        return sums, count * _REDUCER["world"]
    return sums, count


@dataclass
class ConvSpec:
    """kind 'conv' = nn.Conv2d (weight [cout,cin,k,k]); 'convT' = nn.ConvTranspose2d (weight [cin,cout,k,k])."""
    kind: str
    cin: int
    cout: int
    k: int
    stride: int
    pad: int

    @property
    def fwd_mode(self) -> int:
        return MODE_CONV if self.kind == "conv" else MODE_CONVT

    @property
    def bwd_mode(self) -> int:
        return MODE_CONVT if self.kind == "conv" else MODE_CONV

    def out_hw(self, h: int, w: int) -> Tuple[int, int]:
        if self.kind == "conv":
            return (h + 2 * self.pad - self.k) // self.stride + 1, (w + 2 * self.pad - self.k) // self.stride + 1
        return (h - 1) * self.stride - 2 * self.pad + self.k, (w - 1) * self.stride - 2 * self.pad + self.k

    def pack_fwd(self, weight: torch.Tensor) -> torch.Tensor:
        return ops.pack_weight(weight.detach().contiguous(), self.kind == "conv", act_dtype())

    def pack_bwd(self, weight: torch.Tensor) -> torch.Tensor:
        return ops.pack_weight(weight.detach().contiguous(), self.kind != "conv", act_dtype())

    def weight_shape(self) -> Tuple[int, int, int, int]:
        if self.kind == "conv":
            return (self.cout, self.cin, self.k, self.k)
        return (self.cin, self.cout, self.k, self.k)


@dataclass
class Saved:
    """What one conv(+BN+act) layer keeps for backward."""
    x_view: Optional[GsbView] = None
    x_keep: Optional[torch.Tensor] = None     # keeps x's storage alive
    y: Optional[torch.Tensor] = None          # raw conv output (pre-BN), NHWC
    scale: Optional[torch.Tensor] = None
    shift: Optional[torch.Tensor] = None
    mean: Optional[torch.Tensor] = None
    invstd: Optional[torch.Tensor] = None
    count: float = 0.0
    weight: Optional[torch.Tensor] = None     # fp32 weight in torch layout as used in forward
    drop_seed: int = 0
    drop_offset: int = 0                      # global element index of the tensor's first element (Philox dropout)
    drop_mask: Optional[torch.Tensor] = None
    center: Optional[torch.Tensor] = None     # = mean in the centered (fp32) form, else None
    extra: Dict[str, object] = field(default_factory=dict)


def _nhwc_empty(n: int, h: int, w: int, c: int, device, dtype=None, zero: bool = False) -> torch.Tensor:
    f = torch.zeros if zero else torch.empty
    return f((n, h, w, c), device=device, dtype=dtype or act_dtype())


class WeightPacks:
    """Packs the conv weights of a whole network (forward layout and, for training, the transposed backward layout)
    in ONE launch instead of two per layer."""

    def __init__(self):
        self.jobs, self.map = [], {}

    def add(self, key: str, spec: "ConvSpec", weight: torch.Tensor, need_bwd: bool) -> None:
        w = weight.detach().contiguous()
        taps = spec.k * spec.k
        conv = spec.kind == "conv"
        # forward layout [taps, cout, cin]; backward layout [taps, cin, cout]
        wp = torch.empty((taps, spec.cout, spec.cin), device=w.device, dtype=act_dtype())
        self.jobs.append((w, conv, wp))
        wpb = None
        if need_bwd:
            wpb = torch.empty((taps, spec.cin, spec.cout), device=w.device, dtype=act_dtype())
            self.jobs.append((w, not conv, wpb))
        self.map[key] = (wp, wpb)

    def run(self) -> None:
        ops.pack_weight_multi(self.jobs, act_dtype())
        self.jobs = []

    def get(self, key: str):
        return self.map.get(key, (None, None))


_SIDE_STREAMS: Dict[int, "torch.cuda.Stream"] = {}


def pool_fusion_enabled() -> bool:
    """BatchNorm + ReLU + MaxPool2d(2) in one kernel at the encoder's DoubleConv -> Down seam (``GSB_POOL_FUSION=0``: off)."""
    import os
    return os.environ.get("GSB_POOL_FUSION", "1") != "0"


def eval_fold_enabled() -> bool:
    """Inference: BatchNorm (running statistics) + activation folded into the convolution's epilogue (``GSB_EVAL_FOLD=0``: off)."""
    import os
    return os.environ.get("GSB_EVAL_FOLD", "1") != "0"


def wgrad_stream_enabled() -> bool:
    """Weight-gradient kernels on a second stream (default on; ``GSB_WGRAD_STREAM=0`` switches it off): see BackwardBatch.side."""
    import os
    return os.environ.get("GSB_WGRAD_STREAM", "1") != "0"


def _side_stream(device) -> "torch.cuda.Stream":
    idx = device.index if device.index is not None else torch.cuda.current_device()
    st = _SIDE_STREAMS.get(idx)
    if st is None:
        st = _SIDE_STREAMS[idx] = torch.cuda.Stream(device=idx)
    return st


class _OnSide:
    """with batch.on_side(keep): the body's launches go to the batch's side stream, ordered after everything the main
    stream has been given so far; ``keep`` (the tensors the body reads) stays referenced until the next join, so the
    caching allocator cannot hand their memory to a later main-stream allocation while the side stream still reads it."""

    def __init__(self, batch: "BackwardBatch", keep):
        self.batch, self.keep, self.ctx = batch, keep, None

    def __enter__(self):
        b = self.batch
        b.keep.append(self.keep)
        ev = torch.cuda.Event()
        ev.record()
        b.side.wait_event(ev)
        self.ctx = torch.cuda.stream(b.side)
        self.ctx.__enter__()
        b.forked = True
        return b.side

    def __exit__(self, *exc):
        return self.ctx.__exit__(*exc)


class BackwardBatch:
    """Weight gradients of one backward pass: every layer's tensor-core weight gradient arrives as partial-sum slices
    (``ops.conv_wgrad_slices``: plain stores, no atomics), is un-packed -- slices added in slice order -- by ONE batched
    launch per flush into ONE flat gradient buffer laid out in backward order, so that a prefix of that buffer can be
    handed to the gradient all-reduce while the rest of the backward pass is still running.

    ``side``: the weight-gradient kernels of a layer depend on dY only and nothing on the backward chain depends on them
    before the un-pack; they are launched on a second stream (fork after the
    data gradient of the layer, join in ``flush``), so that the tensor-bound weight gradient of layer L runs under the
    HBM-bound BatchNorm / pooling / upsampling backward kernels of layer L-1 (fork / join are events: graph capture
    records them as dependencies)."""

    def __init__(self, device, sizes: Dict[str, int]):
        self.side = _side_stream(device) if (wgrad_stream_enabled() and device.type == "cuda") else None
        self.keep: List[object] = []
        self.forked = False
        total = sum(sizes.values())
        self.gflat = torch.empty(max(total, 1), device=device, dtype=torch.float32)
        self.sizes = dict(sizes)
        self.offsets: Dict[str, Tuple[int, int]] = {}
        self.cursor = 0          # bump allocator: offsets follow the order of first use (= backward order)
        self.flushed = 0
        self.jobs = []
        self.pending = []        # (work handle, tensor) of all-reduces in flight

    def reserve(self, key: str) -> Optional[Tuple[int, int]]:
        """[offset, offset + n) of the flat gradient buffer for this layer (assigned at first use)."""
        if key not in self.sizes:
            return None
        if key not in self.offsets:
            self.offsets[key] = (self.cursor, self.sizes[key])
            self.cursor += self.sizes[key]
        return self.offsets[key]

    def defer_unpack(self, key: str, ws: torch.Tensor, nslices: int, weight_shape, rows_from_d0: bool) -> torch.Tensor:
        off, n = self.reserve(key)
        out = self.gflat[off:off + n].view(weight_shape)
        self.jobs.append((ws, rows_from_d0, out, int(nslices)))
        return out

    def on_side(self, keep) -> _OnSide:
        return _OnSide(self, keep)

    def join(self) -> None:
        """The current stream waits for the side stream's work; the kept inputs may be released after that."""
        if self.forked:
            ev = torch.cuda.Event()
            ev.record(self.side)
            torch.cuda.current_stream().wait_event(ev)
            self.forked = False
        self.keep = []

    def flush(self) -> Tuple[int, int]:
        """Un-pack what has been deferred so far; returns the [start, end) range of the flat gradient buffer that became
        final with this flush."""
        self.join()
        if self.jobs:
            ops.unpack_wgrad_multi(self.jobs)
            self.jobs = []
        rng = (self.flushed, self.cursor)
        self.flushed = self.cursor
        return rng

    # --- gradient all-reduce inside the backward pass (data parallelism) -------------------------------
    def reduce_async(self) -> None:
        """Flush and start the all-reduce of the gradients finished so far on the collective's own stream."""
        start, end = self.flush()
        gfn = _REDUCER.get("grad_fn")
        if gfn is None or end <= start:
            return
        t = self.gflat[start:end]
        self.pending.append((gfn(t), t))
        reserve = int(_REDUCER.get("sm_reserve", 0))
        if reserve > 0:          # persistent kernels launched from now on leave room for the collective's CTAs
            ops.set_sm_reserve(reserve)

    def reduce_finish(self, small: List[torch.Tensor]) -> bool:
        """Flush the rest, all-reduce it together with the small gradients (BatchNorm, biases, stem / head weights) and
        wait for everything in flight.  Returns True when the gradients have been summed over the ranks."""
        gfn = _REDUCER.get("grad_fn")
        start, end = self.flush()
        if gfn is None:
            return False
        if end > start:
            t = self.gflat[start:end]
            self.pending.append((gfn(t), t))
        small = [g for g in small if g is not None]
        flat_small = None
        if small:
            flat_small = torch.cat([g.reshape(-1) for g in small])
            self.pending.append((gfn(flat_small), flat_small))
        for work, _t in self.pending:
            work.wait()
        self.pending = []
        if int(_REDUCER.get("sm_reserve", 0)) > 0:
            ops.set_sm_reserve(0)
        if flat_small is not None:
            torch._foreach_copy_(small, [c.view_as(g) for c, g in zip(flat_small.split([g.numel() for g in small]), small)])
        return True


class AsyncGradReducer:
    """Gradient all-reduce inside a program's backward pass (data parallelism): ``start(t)`` launches the asynchronous
    in-place sum of a finished contiguous gradient on the collective's stream (``parallel._allreduce_async``: NCCL, bf16
    on the wire when configured) while the backward pass goes on; ``finish(small)`` sends the small gradients as one flat
    bucket and waits for everything.  Inactive (``active`` False) without ``overlap_grad_allreduce``."""

    def __init__(self):
        self.gfn = _REDUCER.get("grad_fn") if int(_REDUCER.get("world", 1)) > 1 else None
        self.pending = []
        self.reserved = False

    @property
    def active(self) -> bool:
        return self.gfn is not None

    def start(self, t: torch.Tensor) -> None:
        self.pending.append(self.gfn(t.view(-1)))
        reserve = int(_REDUCER.get("sm_reserve", 0))
        if reserve > 0 and not self.reserved:     # persistent tensor-core kernels leave room for the collective's CTAs
            ops.set_sm_reserve(reserve)
            self.reserved = True

    def wait_all(self) -> None:
        for w in self.pending:
            w.wait()
        self.pending = []

    def finish(self, small: List[torch.Tensor]) -> None:
        small = [g for g in small if g is not None]
        flat = None
        if small:
            flat = torch.cat([g.reshape(-1) for g in small])
            self.pending.append(self.gfn(flat))
        self.wait_all()
        if self.reserved:
            ops.set_sm_reserve(0)
            self.reserved = False
        if flat is not None:
            torch._foreach_copy_(small, [c.view_as(g) for c, g in zip(flat.split([g.numel() for g in small]), small)])


def conv_weight_grad(spec: "ConvSpec", x_view: GsbView, dy_view: GsbView, device, use_tc: int,
                     batch: Optional["BackwardBatch"] = None, key: Optional[str] = None) -> torch.Tensor:
    """Weight gradient of one convolution in torch layout.  The kernel writes partial-sum slices (one per split of the pixel
    range); they are added in slice order while un-packing -- immediately, or by the batch's next flush."""
    shape = (spec.k * spec.k, spec.cout, spec.cin)
    n = ops.conv_wgrad_plan(x_view, dy_view, spec.fwd_mode, spec.k, spec.stride, spec.pad, use_tc=use_tc)
    ws = torch.empty((n,) + shape, device=device, dtype=torch.float32)
    used = ops.conv_wgrad_slices(x_view, dy_view, ws, n, spec.fwd_mode, spec.k, spec.stride, spec.pad, use_tc=use_tc)
    if batch is not None and key is not None and key in batch.sizes:
        return batch.defer_unpack(key, ws, used, spec.weight_shape(), spec.kind == "conv")
    return ops.unpack_wgrad(ws, spec.weight_shape(), spec.kind == "conv", nslices=used)


def _is_stem(s: "ConvSpec", x_view: GsbView, x_keep) -> bool:
    return (s.kind == "conv" and s.cin <= 4 and s.cin * s.k * s.k <= 64 and s.cout % 8 == 0 and s.cout <= 256
            and 256 % (s.cout // 8) == 0 and isinstance(x_keep, torch.Tensor) and x_keep.dim() == 4
            and x_keep.dtype == torch.float32 and x_keep.is_contiguous() and x_keep.shape[1] == s.cin
            and x_view.sw == 1 and x_view.dtype == 0)


class ConvBNAct:
    """conv (+bias) -> [BatchNorm2d] -> act1(dropout(.)) [-> second output act2(.)].

    Covers DoubleConv's halves (unet/unet_parts.py:16-21), the generator's down / up steps
    (networks.py:582-607) and the discriminator blocks (networks.py:640-657).
    """

    def __init__(self, spec: ConvSpec, has_bn: bool, act1: int, act2: Optional[int] = None, dropout: bool = False,
                 instance_norm: bool = False):
        """instance_norm: the normalisation is nn.InstanceNorm2d(affine=False, track_running_stats=False)
        (networks.py:34-35, --norm instance): statistics per image and channel in training AND evaluation, no
        parameters, no buffers.  Same kernels as BatchNorm with a grid dimension over the images."""
        self.spec, self.has_bn, self.act1, self.act2, self.dropout = spec, has_bn, act1, act2, dropout
        self.instance = bool(instance_norm) and has_bn

    # x_view: GsbView of the input (NHWC act buffer or the NCHW fp32 network input)
    def forward(self, x_view: GsbView, x_keep, n: int, h: int, w: int, device, weight, bias, bn, training: bool,
                out1: Optional[torch.Tensor] = None, out2: Optional[torch.Tensor] = None, save: bool = True,
                drop_seed: int = 0, drop_mask: Optional[torch.Tensor] = None, packed=None,
                pool_out: Optional[torch.Tensor] = None):
        """pool_out: optional [n, oh/2, ow/2, cout] tensor that receives MaxPool2d(2) of out1 in the same pass when the
        layer is the plain BatchNorm + ReLU form (``sv.extra['pooled']`` tells the caller whether it was written)."""
        s = self.spec
        oh, ow = s.out_hw(h, w)
        stem = _is_stem(s, x_view, x_keep)
        if (not training and not save and self.has_bn and not self.instance and self.act2 is None and bias is None
                and s.kind == "conv" and _CONFIG["use_tc"] and act_dtype() == torch.bfloat16
                and not centered_bn() and eval_fold_enabled()):
            # inference: BatchNorm of the running statistics folded into the convolution's epilogue -- Conv -> BN -> act is ONE
            # pass, the pre-normalisation tensor is never written (nothing is saved: no backward through this call)
            gamma, beta, rm, rv, nbt = bn
            scale, shift, _m, _i = ops.bn_finalize(None, float(n * oh * ow), gamma, beta, rm, rv, None, BN_MOMENTUM, BN_EPS,
                                                   s.cout, device, centered=False)
            dst = out1 if out1 is not None else _nhwc_empty(n, oh, ow, s.cout, device)
            if stem:
                folded = ops.stem_conv_fwd_affine(x_keep, weight.detach().contiguous(), scale, shift, dst, s.k, s.stride, s.pad,
                                                  self.act1)
            else:
                wp = packed[0] if packed is not None and packed[0] is not None else s.pack_fwd(weight)
                folded = ops.conv_gather_affine(x_view, wp, scale, shift, view_nhwc(dst), s.fwd_mode, s.k, s.stride, s.pad,
                                                self.act1)
            if folded:
                sv = Saved()
                sv.extra["pooled"] = False
                return dst, sv
        y = _nhwc_empty(n, oh, ow, s.cout, device)
        if stem:   # fp32 NCHW network input with 1..4 channels: bandwidth kernel, torch-layout fp32 weights
            stats, rows = ops.stem_conv_fwd(x_keep, weight.detach().contiguous(), bias, y, s.k, s.stride, s.pad,
                                            want_stats=self.has_bn and training and _CONFIG["fused_stats"] and not self.instance)
        else:
            wp = packed[0] if packed is not None and packed[0] is not None else s.pack_fwd(weight)
            want_stats = (self.has_bn and training and _CONFIG["fused_stats"] and not self.instance
                          and not (s.kind == "convT" and s.stride > 1))
            stats, rows = ops.conv_gather(x_view, wp, bias, view_nhwc(y), s.fwd_mode, s.k, s.stride, s.pad,
                                          use_tc=_CONFIG["use_tc"], want_stats=want_stats)
        sv = Saved()
        count = float(n * oh * ow)
        scale = shift = mean = invstd = None
        centered = centered_bn()
        first = first_global_sample(n)
        drop_offset = first * oh * ow * s.cout
        if self.has_bn and self.instance:
            # InstanceNorm2d(affine=False): statistics per (image, channel) -- the BatchNorm kernels with a second grid
            # dimension over the images, so a layer costs the same launches at any batch size
            if out1 is None:
                out1 = _nhwc_empty(n, oh, ow, s.cout, device)
            use_drop = self.dropout and training
            count = float(oh * ow)
            sums = ops.bn_reduce(y, per_image=True)
            scale, shift, mean, invstd = ops.bn_finalize(sums, count, None, None, None, None, None, BN_MOMENTUM, BN_EPS,
                                                         n * s.cout, device, centered=centered)
            center = mean if centered else None
            ops.bn_act(y, scale, shift, self.act1, out1, self.act2 if self.act2 is not None else ACT_NONE,
                       out2 if self.act2 is not None else None, drop_seed if use_drop else 0, drop_mask if use_drop else None,
                       center=center, drop_offset=drop_offset if use_drop else 0, per_image=True)
            if save:
                sv.x_view, sv.x_keep, sv.y = x_view, x_keep, y
                sv.scale, sv.shift, sv.mean, sv.invstd, sv.count = scale, shift, mean, invstd, count
                sv.weight, sv.center = weight, center
                sv.drop_seed = drop_seed if use_drop else 0
                sv.drop_offset = drop_offset if use_drop else 0
                sv.drop_mask = drop_mask if use_drop else None
                sv.extra["instance"] = True
                sv.extra["train_bn"] = False
                sv.extra["stem"] = stem
                sv.extra["wpb"] = packed[1] if packed is not None else None
            return out1, sv
        if self.has_bn:
            gamma, beta, rm, rv, nbt = bn
            fused_ok, peer, pworld = _fused_stats_mode()
            if training and stats is not None and fused_ok:
                count = count * pworld
                scale, shift, mean, invstd = ops.bn_stats_finalize(stats, rows, s.cout, count, gamma, beta, rm, rv, nbt,
                                                                   BN_MOMENTUM, BN_EPS, peer, centered=centered)
            elif training:
                sums = ops.bn_reduce_partials(stats, rows, s.cout) if stats is not None else ops.bn_reduce(y)
                sums, count = _sync_stats(sums, count)
                scale, shift, mean, invstd = ops.bn_finalize(sums, count, gamma, beta, rm, rv, nbt, BN_MOMENTUM, BN_EPS,
                                                             s.cout, device, centered=centered)
            else:
                scale, shift, mean, invstd = ops.bn_finalize(None, count, gamma, beta, rm, rv, None, BN_MOMENTUM,
                                                             BN_EPS, s.cout, device, centered=centered)
        if out1 is None:
            out1 = _nhwc_empty(n, oh, ow, s.cout, device)
        use_drop = self.dropout and training
        center = mean if (centered and self.has_bn) else None
        pooled = (pool_out is not None and pool_fusion_enabled() and self.has_bn and self.act1 == ACT_RELU and self.act2 is None
                  and not use_drop and center is None and oh % 2 == 0 and ow % 2 == 0)
        if pooled:      # BatchNorm + ReLU + MaxPool2d(2) in one pass: the pooling does not read the activation back
            ops.bn_act_pool(y, scale, shift, out1, pool_out)
        else:
            ops.bn_act(y, scale, shift, self.act1, out1, self.act2 if self.act2 is not None else ACT_NONE,
                       out2 if self.act2 is not None else None, drop_seed if use_drop else 0,
                       drop_mask if use_drop else None, center=center, drop_offset=drop_offset if use_drop else 0)
        sv.extra["pooled"] = pooled
        if save:
            sv.x_view, sv.x_keep, sv.y = x_view, x_keep, y
            sv.scale, sv.shift, sv.mean, sv.invstd, sv.count = scale, shift, mean, invstd, count
            sv.weight = weight
            sv.center = center
            sv.drop_offset = drop_offset if use_drop else 0
            sv.drop_seed = drop_seed if use_drop else 0
            sv.drop_mask = drop_mask if use_drop else None
            sv.extra["train_bn"] = self.has_bn and training
            sv.extra["stem"] = stem
            sv.extra["wpb"] = packed[1] if packed is not None else None
        return out1, sv

    def backward(self, sv: Saved, dA1: torch.Tensor, dA2: Optional[torch.Tensor], need_dx: bool, need_dw: bool,
                 need_bn: bool, need_db: bool, dx_out: Optional[torch.Tensor] = None, dx_view: Optional[GsbView] = None,
                 batch: Optional["BackwardBatch"] = None, key: Optional[str] = None, pool=None):
        """Returns (dx, dweight, dbias, dgamma, dbeta).  dx is written to dx_out / dx_view when given.
        pool = (a, dpooled): this layer's activation ``a`` feeds a MaxPool2d(2) whose gradient ``dpooled`` still has to be
        routed into dA1 (in place): folded into the statistics pass where the layer has the plain BatchNorm + ReLU form,
        else done by the separate max-pool backward kernel first."""
        s = self.spec
        y = sv.y
        n, oh, ow, _ = y.shape
        device = y.device
        train_bn = bool(sv.extra.get("train_bn"))
        pool_grad = None
        if pool is not None:
            fusable = (pool_fusion_enabled() and train_bn and _fused_stats_mode()[0] and sv.extra.get("instance") is None
                       and self.act1 == ACT_RELU and dA2 is None and not sv.drop_seed and sv.drop_mask is None
                       and sv.center is None and y.dtype == torch.bfloat16 and oh % 2 == 0 and ow % 2 == 0
                       and 256 % (s.cout // 8) == 0)
            if fusable:
                pool_grad = pool[1]
            else:
                ops.maxpool2_bwd(pool[0], pool[1], dA1, accumulate=True)
        dY = torch.empty_like(y)
        dgamma = dbeta = None
        act2 = self.act2 if self.act2 is not None else ACT_NONE
        sums = None
        mean = invstd = None
        fused_ok, peer, _pw = _fused_stats_mode()
        inst = sv.extra.get("instance")
        if inst is not None:
            sums, _, _ = ops.bn_bwd_stats(y, dA1, dA2, sv.scale, sv.shift, sv.mean, sv.invstd, self.act1, act2, sv.drop_seed,
                                          sv.drop_mask, False, None, center=sv.center, drop_offset=sv.drop_offset, per_image=True)
            ops.bn_bwd_apply(y, dA1, dA2, sv.scale, sv.shift, sv.mean, sv.invstd, self.act1, act2, sums, sv.count, dY, None, None,
                             sv.drop_seed, sv.drop_mask, center=sv.center, drop_offset=sv.drop_offset, per_image=True)
        elif train_bn and fused_ok:
            sums, dgamma, dbeta = ops.bn_bwd_stats(y, dA1, dA2, sv.scale, sv.shift, sv.mean, sv.invstd, self.act1, act2,
                                                   sv.drop_seed, sv.drop_mask, need_bn, peer, center=sv.center,
                                                   drop_offset=sv.drop_offset, pool_grad=pool_grad)
            mean, invstd = sv.mean, sv.invstd
        elif train_bn:
            sums = ops.bn_bwd_reduce(y, dA1, dA2, sv.scale, sv.shift, sv.mean, sv.invstd, self.act1, act2,
                                     sv.drop_seed, sv.drop_mask, center=sv.center, drop_offset=sv.drop_offset)
            if need_bn:   # parameter gradients from the LOCAL sums (the gradient all-reduce adds the other ranks)
                dgamma, dbeta = ops.bn_param_grad(sums, s.cout)
            sums, _ = _sync_stats(sums, 0.0)
            mean, invstd = sv.mean, sv.invstd
        elif self.has_bn and need_bn:
            # eval-mode BatchNorm (running statistics are constants): dgamma = sum dy*xhat_running, dbeta = sum dy
            es = ops.bn_bwd_reduce(y, dA1, dA2, sv.scale, sv.shift, sv.mean, sv.invstd, self.act1, act2,
                                   sv.drop_seed, sv.drop_mask, center=sv.center, drop_offset=sv.drop_offset)
            dgamma, dbeta = ops.bn_param_grad(es, s.cout)
        if inst is None:
            ops.bn_bwd_apply(y, dA1, dA2, sv.scale, sv.shift, mean, invstd, self.act1, act2, sums, sv.count, dY,
                             None, None, sv.drop_seed, sv.drop_mask, center=sv.center, drop_offset=sv.drop_offset)
        dyv = view_nhwc(dY)
        dweight = dbias = None
        side_wgrad = False
        if need_dw and sv.extra.get("stem"):
            dweight = ops.stem_conv_wgrad(sv.x_keep, dY, s.k, s.stride, s.pad)
        elif need_dw and batch is not None and batch.side is not None and key is not None and key in batch.sizes:
            side_wgrad = True                   # launched AFTER the data gradient, on the side stream (below)
        elif need_dw:
            dweight = conv_weight_grad(s, sv.x_view, dyv, device, _CONFIG["use_tc"], batch, key)
        if need_db:
            dbias = torch.empty(s.cout, device=device, dtype=torch.float32)
            ops.bias_grad(dyv, dbias)
        dx = None
        if need_dx:
            wpb = sv.extra.get("wpb")
            if wpb is None:
                wpb = s.pack_bwd(sv.weight)
            if dx_view is None:
                if dx_out is None:
                    xv = sv.x_view
                    dx_out = _nhwc_empty(xv.n, xv.h, xv.w, xv.c, device)
                dx_view = view_nhwc(dx_out)
            ops.conv_gather(dyv, wpb, None, dx_view, s.bwd_mode, s.k, s.stride, s.pad, use_tc=_CONFIG["use_tc"])
            dx = dx_out
        if side_wgrad:
            # fork AFTER the data gradient: queued behind it, the weight gradient starts when the data gradient's CTAs
            # drain and runs under the HBM-bound BatchNorm / pooling / upsampling backward kernels of the next layer
            with batch.on_side((dY, sv)):
                dweight = conv_weight_grad(s, sv.x_view, dyv, device, _CONFIG["use_tc"], batch, key)
        return dx, dweight, dbias, dgamma, dbeta


def _edge_tc() -> int:
    """use_tc of the network-edge convolutions: 2 (tensor cores where a kernel exists for the shape -- the generator's CW -> 1
    transposed 8x8 cell, thin_tc.cu -- else the thin SIMT kernels) unless the tensor-core mode is switched off."""
    return 2 if _CONFIG["use_tc"] else 0


class EdgeConv:
    """conv (+bias) -> act, written straight to an NCHW fp32 tensor (network outputs with 1..3
    channels: OutConv unet_parts.py:71-77, the discriminator's last conv networks.py:660, the
    generator's outermost up-cell + Tanh networks.py:590-592)."""

    def __init__(self, spec: ConvSpec, act: int = ACT_NONE):
        self.spec, self.act = spec, act

    def forward(self, x: torch.Tensor, weight, bias, save: bool = True):
        s = self.spec
        n, h, w, _ = x.shape
        oh, ow = s.out_hw(h, w)
        head = self._is_head(x)
        if head:
            out = ops.head_conv_fwd(x, weight.detach().contiguous(), bias)
        else:
            out = torch.empty((n, s.cout, oh, ow), device=x.device, dtype=torch.float32)
            wp = s.pack_fwd(weight)
            ops.conv_gather(view_nhwc(x), wp, bias, view_nchw(out), s.fwd_mode, s.k, s.stride, s.pad, act=self.act,
                            use_tc=_edge_tc())
        sv = Saved()
        if save:
            sv.x_view, sv.x_keep, sv.weight = view_nhwc(x), x, weight
            sv.extra["out"] = out if self.act != ACT_NONE else None
            sv.extra["head"] = head
        return out, sv

    def _is_head(self, x: torch.Tensor) -> bool:
        s = self.spec
        return (s.kind == "conv" and s.k == 1 and s.stride == 1 and s.pad == 0 and self.act == ACT_NONE
                and s.cout <= 8 and s.cin % 64 == 0 and s.cin <= 256 and x.stride(3) == 1
                and x.stride(2) % 8 == 0 and x.stride(1) == x.shape[2] * x.stride(2))

    def backward(self, sv: Saved, dout: torch.Tensor, need_dx: bool, need_dw: bool, need_db: bool,
                 dx_out: Optional[torch.Tensor] = None):
        s = self.spec
        dout = dout.contiguous()
        if dout.dtype != torch.float32:
            raise RuntimeError("genseg_b200: gradient of an fp32 output must be fp32")
        if self.act != ACT_NONE:
            dout = ops.act_bwd(sv.extra["out"], dout, self.act)
        dv = view_nchw(dout)
        device = dout.device
        dweight = dbias = dx = None
        if sv.extra.get("head") and dx_out is None:
            return ops.head_conv_bwd(sv.x_keep, sv.weight.detach().contiguous(), dout, need_dx, need_dw, need_db)
        if need_dw:
            dweight = conv_weight_grad(s, sv.x_view, dv, device, _edge_tc())
        if need_db:
            dbias = torch.empty(s.cout, device=device, dtype=torch.float32)
            ops.bias_grad(dv, dbias)
        if need_dx:
            xv = sv.x_view
            if dx_out is None:
                dx_out = _nhwc_empty(xv.n, xv.h, xv.w, xv.c, device)
            wpb = s.pack_bwd(sv.weight)
            ops.conv_gather(dv, wpb, None, view_nhwc(dx_out), s.bwd_mode, s.k, s.stride, s.pad, use_tc=_edge_tc())
            dx = dx_out
        return dx, dweight, dbias
