"""Stand-alone programs for the U-Net parts (DoubleConv / Down / Up / OutConv used on their own).

Same kernels as the whole-network schedule; the NCHW fp32 API edge is converted to/from the NHWC
compute layout with gsb_copy_view.  The whole-network path (unet_program.py) is the tuned one.
"""
from __future__ import annotations

from typing import Dict

import torch

from . import ops
from ._lib import view_nchw, view_nhwc
from .engine import ConvSpec, EdgeConv, _nhwc_empty, act_dtype
from .unet_program import DoubleConvOps, _put


def _f32c(x: torch.Tensor) -> torch.Tensor:
    x = x.contiguous()
    return x if x.dtype == torch.float32 else x.float()


class DoubleConvProgram:
    def __init__(self, cin: int, cout: int, mid: int, prefix: str = ""):
        self.dc = DoubleConvOps(prefix.rstrip("."), cin, cout, mid) if prefix else _Rooted(cin, cout, mid)
        p = self.dc.prefix + ".double_conv" if self.dc.prefix else "double_conv"
        self.param_names = [p + ".0.weight", p + ".1.weight", p + ".1.bias", p + ".3.weight", p + ".4.weight",
                            p + ".4.bias"]

    def forward(self, x, P, B, training, save, extra=None):
        x = _f32c(x)
        n, _, h, w = x.shape
        a, sv = self.dc.forward(view_nchw(x), x, n, h, w, x.device, P, B, training, save)
        return ops.nhwc_to_nchw(a), {"dc": sv, "x": x}

    def backward(self, tape, dout, need, need_dx):
        G: Dict[str, torch.Tensor] = {}
        d = ops.nchw_to_nhwc(_f32c(dout), act_dtype())
        dx = torch.empty_like(tape["x"]) if need_dx else None
        self.dc.backward(tape["dc"], d, G, need, need_dx, dx_out=dx, dx_view=view_nchw(dx) if need_dx else None)
        return dx, G


class _Rooted(DoubleConvOps):
    """DoubleConvOps whose parameter names have no module prefix ('double_conv.0.weight', ...)."""

    def __init__(self, cin, cout, mid):
        super().__init__("", cin, cout, mid)

    def _bn(self, P, B, idx):
        p = f"double_conv.{idx}"
        return (P[p + ".weight"], P[p + ".bias"], B[p + ".running_mean"], B[p + ".running_var"],
                B[p + ".num_batches_tracked"])

    def forward(self, x_view, x_keep, n, h, w, device, P, B, training, save, out=None):
        self.prefix = ""
        return _prefixed(super().forward, self, x_view, x_keep, n, h, w, device, P, B, training, save, out)

    def backward(self, saved, dA, G, need, need_dx, dx_out=None, dx_view=None):
        P2 = _Shift(need)
        G2: Dict[str, torch.Tensor] = {}
        dx = super().backward(saved, dA, G2, P2, need_dx, dx_out=dx_out, dx_view=dx_view)
        for k, v in G2.items():
            G[k[1:]] = v
        return dx


class _Shift(dict):
    """Maps '.double_conv.X' lookups (empty prefix + '.') onto 'double_conv.X' keys."""

    def __init__(self, base):
        super().__init__()
        self.base = base

    def __getitem__(self, k):
        return self.base[k[1:] if k.startswith(".") else k]

    def get(self, k, default=None):
        return self.base.get(k[1:] if k.startswith(".") else k, default)


def _prefixed(fn, self, x_view, x_keep, n, h, w, device, P, B, training, save, out):
    return fn(x_view, x_keep, n, h, w, device, _Shift(P), _Shift(B), training, save, out=out)


class DownProgram:
    """MaxPool2d(2) -> DoubleConv, unet_parts.py:28-39."""

    def __init__(self, cin: int, cout: int):
        self.dc = DoubleConvOps("maxpool_conv.1", cin, cout)
        p = "maxpool_conv.1.double_conv"
        self.param_names = [p + ".0.weight", p + ".1.weight", p + ".1.bias", p + ".3.weight", p + ".4.weight",
                            p + ".4.bias"]

    def forward(self, x, P, B, training, save, extra=None):
        x = _f32c(x)
        xh = ops.nchw_to_nhwc(x, act_dtype())
        pooled = ops.maxpool2_fwd(xh)
        n, h, w, _ = pooled.shape
        a, sv = self.dc.forward(view_nhwc(pooled), pooled, n, h, w, x.device, P, B, training, save)
        return ops.nhwc_to_nchw(a), {"dc": sv, "xh": xh}

    def backward(self, tape, dout, need, need_dx):
        G: Dict[str, torch.Tensor] = {}
        d = ops.nchw_to_nhwc(_f32c(dout), act_dtype())
        dp = self.dc.backward(tape["dc"], d, G, need, need_dx)
        dx = None
        if need_dx:
            dxh = torch.empty_like(tape["xh"])
            ops.maxpool2_bwd(tape["xh"], dp, dxh, accumulate=False)
            dx = ops.nhwc_to_nchw(dxh)
        return dx, G


class UpProgram:
    """Up(x1, x2), unet_parts.py:42-68."""

    def __init__(self, cin: int, cout: int, bilinear: bool):
        from .engine import ConvBNAct
        from ._lib import ACT_NONE
        self.bilinear = bilinear
        self.cin = cin
        self.dc = DoubleConvOps("conv", cin, cout, cin // 2 if bilinear else None)
        p = "conv.double_conv"
        self.param_names = ([] if bilinear else ["up.weight", "up.bias"]) + \
            [p + ".0.weight", p + ".1.weight", p + ".1.bias", p + ".3.weight", p + ".4.weight", p + ".4.bias"]
        self.upconv = None if bilinear else ConvBNAct(ConvSpec("convT", cin, cin // 2, 2, 2, 0), False, ACT_NONE)

    def forward(self, xs, P, B, training, save, extra=None):
        x1, x2 = _f32c(xs[0]), _f32c(xs[1])
        dev = x1.device
        n, c2, hh, ww = x2.shape
        a1 = ops.nchw_to_nhwc(x1, act_dtype())
        h1, w1 = a1.shape[1], a1.shape[2]
        c_up = a1.shape[3] if self.bilinear else self.cin // 2
        cat = _nhwc_empty(n, hh, ww, c2 + c_up, dev, zero=True)
        s = torch.cuda.current_stream(dev).cuda_stream
        ops.copy_view(view_nchw(x2), view_nhwc(cat[..., :c2]), s)
        pad_t, pad_l = (hh - 2 * h1) // 2, (ww - 2 * w1) // 2
        upper = cat[..., c2:]
        tape = {"pad": (pad_t, pad_l), "c2": c2, "a1": a1, "shape1": tuple(x1.shape), "shape2": tuple(x2.shape)}
        if self.bilinear:
            ops.upsample2_fwd(a1, upper, pad_t, pad_l)
        else:
            o, sv = self.upconv.forward(view_nhwc(a1), a1, n, h1, w1, dev, P["up.weight"], P["up.bias"], None, training,
                                        save=save)
            ops.copy_view(view_nhwc(o), view_nhwc(upper[:, pad_t:pad_t + 2 * h1, pad_l:pad_l + 2 * w1, :]), s)
            tape["up"] = sv
        a, sv = self.dc.forward(view_nhwc(cat), cat, n, hh, ww, dev, P, B, training, save)
        tape["dc"] = sv
        return ops.nhwc_to_nchw(a), tape

    def backward(self, tape, dout, need, need_dx):
        G: Dict[str, torch.Tensor] = {}
        d = ops.nchw_to_nhwc(_f32c(dout), act_dtype())
        n, hh, ww, _ = d.shape
        dev = d.device
        c2 = tape["c2"]
        c_cat = tape["dc"][0].x_view.c
        dcat = _nhwc_empty(n, hh, ww, c_cat, dev)
        self.dc.backward(tape["dc"], d, G, need, True, dx_out=dcat)
        pad_t, pad_l = tape["pad"]
        a1 = tape["a1"]
        h1, w1 = a1.shape[1], a1.shape[2]
        s = torch.cuda.current_stream(dev).cuda_stream
        dx2 = torch.empty(tape["shape2"], device=dev, dtype=torch.float32)
        ops.copy_view(view_nhwc(dcat[..., :c2]), view_nchw(dx2), s)
        dupper = dcat[..., c2:]
        if self.bilinear:
            d1 = torch.empty_like(a1)
            ops.upsample2_bwd(dupper, d1, pad_t, pad_l)
        else:
            win = _nhwc_empty(n, 2 * h1, 2 * w1, dupper.shape[3], dev)
            ops.copy_view(view_nhwc(dupper[:, pad_t:pad_t + 2 * h1, pad_l:pad_l + 2 * w1, :]), view_nhwc(win), s)
            d1, dw, db, _, _ = self.upconv.backward(tape["up"], win, None, True, need.get("up.weight", False), False,
                                                    need.get("up.bias", False))
            _put(G, "up.weight", dw)
            _put(G, "up.bias", db)
        dx1 = ops.nhwc_to_nchw(d1)
        return (dx1, dx2), G


class OutConvProgram:
    def __init__(self, cin: int, cout: int):
        self.conv = EdgeConv(ConvSpec("conv", cin, cout, 1, 1, 0))
        self.param_names = ["conv.weight", "conv.bias"]

    def forward(self, x, P, B, training, save, extra=None):
        xh = ops.nchw_to_nhwc(_f32c(x), act_dtype())
        out, sv = self.conv.forward(xh, P["conv.weight"], P["conv.bias"], save)
        return out, {"sv": sv}

    def backward(self, tape, dout, need, need_dx):
        G: Dict[str, torch.Tensor] = {}
        dx, dw, db = self.conv.backward(tape["sv"], dout, need_dx, need.get("conv.weight", False),
                                        need.get("conv.bias", False))
        _put(G, "conv.weight", dw)
        _put(G, "conv.bias", db)
        return (ops.nhwc_to_nchw(dx) if dx is not None else None), G
