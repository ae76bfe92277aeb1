"""Whole-network forward/backward schedule of the 2-D U-Net (unet/unet_model.py:26-37).

Buffer plan (NHWC, bf16 in production):
  * every skip tensor x1..x4 is written by its producer (BN+ReLU pass) directly into the LOWER
    channel half of the decoder's concat buffer; the bilinear upsample (or ConvTranspose2d k2 s2)
    writes the UPPER half, so ``torch.cat`` (unet_parts.py:67) never materialises;
  * backward mirrors it: the dgrad of the first decoder conv writes the full concat gradient, the
    upsample backward reads its upper half, the max-pool backward accumulates into the lower half.
"""
from __future__ import annotations

from typing import Dict, List, Optional, Sequence, Tuple

import torch

from . import ops
from ._lib import ACT_NONE, ACT_RELU, view_nchw, view_nhwc
from .engine import BackwardBatch, ConvBNAct, ConvSpec, EdgeConv, Saved, WeightPacks, _nhwc_empty, act_dtype


def unet_param_names(n_channels: int, n_classes: int, bilinear: bool) -> List[str]:
    """Learnable parameters (state_dict names) in the order the program consumes them."""
    names: List[str] = []

    def dc(prefix: str) -> None:
        names.extend([prefix + ".double_conv.0.weight", prefix + ".double_conv.1.weight", prefix + ".double_conv.1.bias",
                      prefix + ".double_conv.3.weight", prefix + ".double_conv.4.weight", prefix + ".double_conv.4.bias"])

    dc("inc")
    for i in range(1, 5):
        dc(f"down{i}.maxpool_conv.1")
    for i in range(1, 5):
        if not bilinear:
            names.extend([f"up{i}.up.weight", f"up{i}.up.bias"])
        dc(f"up{i}.conv")
    names.extend(["outc.conv.weight", "outc.conv.bias"])
    return names


class DoubleConvOps:
    """(conv3x3 -> BN -> ReLU) x 2, unet/unet_parts.py:8-25."""

    def __init__(self, prefix: str, cin: int, cout: int, mid: Optional[int] = None):
        mid = mid or cout
        self.prefix = prefix
        self.cin, self.mid, self.cout = cin, mid, cout
        self.c0 = ConvBNAct(ConvSpec("conv", cin, mid, 3, 1, 1), True, ACT_RELU)
        self.c1 = ConvBNAct(ConvSpec("conv", mid, cout, 3, 1, 1), True, ACT_RELU)

    def _bn(self, P, B, idx: int):
        p = f"{self.prefix}.double_conv.{idx}"
        return (P[p + ".weight"], P[p + ".bias"], B[p + ".running_mean"], B[p + ".running_var"],
                B[p + ".num_batches_tracked"])

    def conv_layers(self):
        p = self.prefix + ".double_conv"
        return [(p + ".0.weight", self.c0.spec), (p + ".3.weight", self.c1.spec)]

    def forward(self, x_view, x_keep, n, h, w, device, P, B, training, save, out=None, packs=None, pool_out=None):
        p = self.prefix + ".double_conv"
        pk0 = packs.get(p + ".0.weight") if packs is not None else None
        pk1 = packs.get(p + ".3.weight") if packs is not None else None
        a0, s0 = self.c0.forward(x_view, x_keep, n, h, w, device, P[p + ".0.weight"], None, self._bn(P, B, 1),
                                 training, save=save, packed=pk0)
        a1, s1 = self.c1.forward(view_nhwc(a0), a0, n, h, w, device, P[p + ".3.weight"], None, self._bn(P, B, 4),
                                 training, out1=out, save=save, packed=pk1, pool_out=pool_out)
        return a1, (s0, s1)

    def backward(self, saved, dA, G: Dict[str, torch.Tensor], need: Dict[str, bool], need_dx: bool,
                 dx_out=None, dx_view=None, batch=None, pool=None):
        s0, s1 = saved
        p = self.prefix + ".double_conv"
        nw1, nb1 = need.get(p + ".3.weight", False), need.get(p + ".4.weight", False) or need.get(p + ".4.bias", False)
        d0, dw, _, dg, db = self.c1.backward(s1, dA, None, True, nw1, nb1, False, batch=batch, key=p + ".3.weight", pool=pool)
        _put(G, p + ".3.weight", dw)
        _put(G, p + ".4.weight", dg)
        _put(G, p + ".4.bias", db)
        nw0, nb0 = need.get(p + ".0.weight", False), need.get(p + ".1.weight", False) or need.get(p + ".1.bias", False)
        dx, dw, _, dg, db = self.c0.backward(s0, d0, None, need_dx, nw0, nb0, False, dx_out=dx_out, dx_view=dx_view,
                                             batch=batch, key=p + ".0.weight")
        _put(G, p + ".0.weight", dw)
        _put(G, p + ".1.weight", dg)
        _put(G, p + ".1.bias", db)
        return dx


def _put(G: Dict[str, torch.Tensor], k: str, v: Optional[torch.Tensor]) -> None:
    if v is not None:
        G[k] = v


class UNetProgram:
    def __init__(self, n_channels: int, n_classes: int, bilinear: bool):
        self.n_channels, self.n_classes, self.bilinear = n_channels, n_classes, bilinear
        f = 2 if bilinear else 1
        self.inc = DoubleConvOps("inc", n_channels, 64)
        chans = [64, 128, 256, 512, 1024 // f]
        self.chans = chans
        self.downs = [DoubleConvOps(f"down{i}.maxpool_conv.1", chans[i - 1], chans[i]) for i in range(1, 5)]
        ups_io = [(1024, 512 // f), (512, 256 // f), (256, 128 // f), (128, 64)]
        self.ups = []
        self.upconvs = []
        for i, (cin, cout) in enumerate(ups_io, start=1):
            if bilinear:
                self.ups.append(DoubleConvOps(f"up{i}.conv", cin, cout, cin // 2))
                self.upconvs.append(None)
            else:
                self.ups.append(DoubleConvOps(f"up{i}.conv", cin, cout))
                # nn.ConvTranspose2d(cin, cin//2, 2, 2) with bias, unet_parts.py:53
                self.upconvs.append(ConvBNAct(ConvSpec("convT", cin, cin // 2, 2, 2, 0), False, ACT_NONE))
        self.outc = EdgeConv(ConvSpec("conv", 64, n_classes, 1, 1, 0))
        self.param_names = unet_param_names(n_channels, n_classes, bilinear)

    # ------------------------------------------------------------------------------------------
    def forward(self, x: torch.Tensor, P: Dict[str, torch.Tensor], B: Dict[str, torch.Tensor], training: bool,
                save: bool, extra=None):
        if x.dim() != 4 or x.shape[1] != self.n_channels:
            raise RuntimeError("UNet: expected input [N,%d,H,W], got %s" % (self.n_channels, tuple(x.shape)))
        x = x.contiguous()
        if x.dtype != torch.float32:
            x = x.float()
        n, _, h, w = x.shape
        if h < 16 or w < 16:
            raise RuntimeError("UNet: spatial size must be >= 16 (four 2x2 poolings)")
        dev = x.device
        tape: Dict[str, object] = {"x": x}
        sizes = [(h, w)]
        for _ in range(4):
            sizes.append((sizes[-1][0] // 2, sizes[-1][1] // 2))
        # concat buffers for the four decoder levels: cat[k] joins skip x_{k+1} (lower half) with the upsampled path
        cats = []
        for k in range(4):
            ck = self.chans[k]
            up_c = ck if self.bilinear else ck  # upper half has as many channels as the skip in both variants
            hh, ww = sizes[k]
            need_zero = (2 * sizes[k + 1][0] != hh) or (2 * sizes[k + 1][1] != ww)
            cats.append(_nhwc_empty(n, hh, ww, ck + up_c, dev, zero=need_zero and not self.bilinear))
        # every tensor-core conv weight packed (forward + backward layouts) in one launch
        packs = WeightPacks()
        stem_key = "inc.double_conv.0.weight" if self.n_channels <= 4 else None
        for blk in [self.inc] + self.downs + self.ups:
            for key, spec in blk.conv_layers():
                if key != stem_key:
                    packs.add(key, spec, P[key], need_bwd=save)
        packs.run()
        # encoder
        # the block in front of a Down writes its max-pooled activation in the same pass as the activation itself
        def pool_buffer(level):
            return _nhwc_empty(n, sizes[level + 1][0], sizes[level + 1][1], self.chans[level], dev)

        pooled = pool_buffer(0)
        x1, sv = self.inc.forward(view_nchw(x), x, n, h, w, dev, P, B, training, save, out=cats[0][..., :64], packs=packs,
                                  pool_out=pooled)
        tape["inc"] = sv
        skips = [x1]
        cur = x1
        for i in range(4):
            if not sv[1].extra.get("pooled"):           # odd size, fp32 validation mode, ...: separate pooling pass
                pooled = ops.maxpool2_fwd(cur)
            hh, ww = sizes[i + 1]
            out = cats[i + 1][..., :self.chans[i + 1]] if i < 3 else None
            nxt = pool_buffer(i + 1) if i < 3 else None
            cur, sv = self.downs[i].forward(view_nhwc(pooled), pooled, n, hh, ww, dev, P, B, training, save, out=out,
                                            packs=packs, pool_out=nxt)
            pooled = nxt
            tape[f"down{i + 1}"] = sv
            skips.append(cur)
        tape["skips"] = skips if save else None
        # decoder
        for i in range(4):
            lvl = 3 - i                       # concat level joined here
            cat = cats[lvl]
            c_skip = self.chans[lvl]
            hh, ww = sizes[lvl]
            sh_, sw_ = cur.shape[1], cur.shape[2]
            pad_t, pad_l = (hh - 2 * sh_) // 2, (ww - 2 * sw_) // 2
            upper = cat[..., c_skip:]
            if self.bilinear:
                ops.upsample2_fwd(cur, upper, pad_t, pad_l)
                tape[f"up{i + 1}.in_shape"] = tuple(cur.shape)
            else:
                name = f"up{i + 1}.up"
                full = (2 * sh_ == hh) and (2 * sw_ == ww)
                window = upper[:, pad_t:pad_t + 2 * sh_, pad_l:pad_l + 2 * sw_, :]
                o1, sv = self.upconvs[i].forward(view_nhwc(cur), cur, n, sh_, sw_, dev, P[name + ".weight"],
                                                 P[name + ".bias"], None, training, out1=window if full else None,
                                                 save=save)
                if not full:  # F.pad border stays zero (buffer was zero-filled)
                    ops.copy_view(view_nhwc(o1), view_nhwc(window), torch.cuda.current_stream(dev).cuda_stream)
                tape[name] = sv
            tape[f"up{i + 1}.pad"] = (pad_t, pad_l)
            cur, sv = self.ups[i].forward(view_nhwc(cat), cat, n, hh, ww, dev, P, B, training, save, packs=packs)
            tape[f"up{i + 1}"] = sv
        logits, sv = self.outc.forward(cur, P["outc.conv.weight"], P["outc.conv.bias"], save)
        tape["outc"] = sv
        tape["sizes"] = sizes
        return logits, tape

    # ------------------------------------------------------------------------------------------
    def backward(self, tape, dlogits: torch.Tensor, need: Dict[str, bool], need_dx: bool):
        G: Dict[str, torch.Tensor] = {}
        sizes = tape["sizes"]
        skips = tape["skips"]
        n = dlogits.shape[0]
        dev = dlogits.device
        sizes_w = {}
        stem_key = "inc.double_conv.0.weight" if self.n_channels <= 4 else None
        for blk in [self.inc] + self.downs + self.ups:
            for key, spec in blk.conv_layers():
                if key != stem_key and need.get(key, False):
                    sizes_w[key] = spec.k * spec.k * spec.cout * spec.cin
        batch = BackwardBatch(dev, sizes_w)
        dcur, dw, db = self.outc.backward(tape["outc"], dlogits, True, need.get("outc.conv.weight", False),
                                          need.get("outc.conv.bias", False))
        _put(G, "outc.conv.weight", dw)
        _put(G, "outc.conv.bias", db)
        dskip: List[Optional[torch.Tensor]] = [None] * 4
        for i in range(3, -1, -1):
            lvl = 3 - i
            c_skip = self.chans[lvl]
            hh, ww = sizes[lvl]
            c_cat = tape[f"up{i + 1}"][0].x_view.c
            dcat = _nhwc_empty(n, hh, ww, c_cat, dev)
            self.ups[i].backward(tape[f"up{i + 1}"], dcur, G, need, True, dx_out=dcat, batch=batch)
            dskip[lvl] = dcat[..., :c_skip]
            dupper = dcat[..., c_skip:]
            pad_t, pad_l = tape[f"up{i + 1}.pad"]
            if self.bilinear:
                shp = tape[f"up{i + 1}.in_shape"]
                dprev = torch.empty(shp, device=dev, dtype=act_dtype())
                ops.upsample2_bwd(dupper, dprev, pad_t, pad_l)
            else:
                name = f"up{i + 1}.up"
                sv: Saved = tape[name]
                sh_, sw_ = sv.x_view.h, sv.x_view.w
                window = dupper[:, pad_t:pad_t + 2 * sh_, pad_l:pad_l + 2 * sw_, :]
                if not ((2 * sh_ == hh) and (2 * sw_ == ww)):
                    tmp = _nhwc_empty(n, 2 * sh_, 2 * sw_, window.shape[3], dev)
                    ops.copy_view(view_nhwc(window), view_nhwc(tmp), torch.cuda.current_stream(dev).cuda_stream)
                    window = tmp
                dprev, dw, db, _, _ = self.upconvs[i].backward(sv, window, None, True, need.get(name + ".weight", False),
                                                               False, need.get(name + ".bias", False))
                _put(G, name + ".weight", dw)
                _put(G, name + ".bias", db)
            dcur = dprev
        # encoder: dcur = d x5
        pool = None
        for i in range(3, -1, -1):
            dpooled = self.downs[i].backward(tape[f"down{i + 1}"], dcur, G, need, True, batch=batch, pool=pool)
            # the pooled gradient still has to go through the 2x2 max-pool into the skip gradient that already holds the
            # concat part: the next block's BatchNorm backward statistics pass does it on the way (engine.ConvBNAct.backward)
            pool = (skips[i], dpooled)
            dcur = dskip[i]
            if i >= 2:
                # i == 3: decoder + bottleneck weight gradients (72 % of the parameters) are final: start their all-reduce
                # now, it runs on the collective's stream under the remaining (most expensive) part of the backward pass;
                # i == 2: down3 (20 %) follows, so that only 7 % of the bytes are left for the exposed tail bucket
                batch.reduce_async()
        dx = None
        dxv = None
        if need_dx:
            dx = torch.empty_like(tape["x"])
            dxv = view_nchw(dx)
        self.inc.backward(tape["inc"], dcur, G, need, need_dx, dx_view=dxv, dx_out=dx, batch=batch, pool=pool)
        gbase = batch.gflat.untyped_storage().data_ptr()
        small = [g for g in G.values() if g is not None and g.untyped_storage().data_ptr() != gbase]
        if batch.reduce_finish(small):
            G["__reduced__"] = True     # functions.ProgramFunction marks the parameters; parallel.allreduce_gradients skips them
        return dx, G
