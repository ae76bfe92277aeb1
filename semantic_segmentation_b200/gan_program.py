"""Forward/backward schedules of the Pix2Pix generator (unet_256 with mixed transposed-conv cells,
models_pix2pix/networks.py:476-617) and the 70x70 PatchGAN discriminator (networks.py:620-665).

Generator notes
  * every up-convolution is the reference's softmax(alpha)-weighted sum of ConvTranspose2d k4/k6/k8 (s2,
    p=(k-2)/2); the three kernels are concentric, so they are merged into ONE 8x8/s2/p3 transposed conv each
    forward (gsb_merge_upconv) and the gradient is un-merged into dW4/dW6/dW8/dalpha;
  * the in-place LeakyReLU of networks.py:583 makes every skip carry lrelu(e); the parent's ReLU over the
    concat then yields relu(e).  One BN pass therefore writes lrelu(BN(y)) (next down conv) AND relu(BN(y))
    (lower half of the concat buffer); the up path writes relu(dropout(BN(y))) into the upper half.
"""
from __future__ import annotations

from typing import Dict, List, Optional

import torch

from . import ops
from ._lib import ACT_LRELU, ACT_NONE, ACT_RELU, ACT_TANH, view_nchw, view_nhwc
from .engine import AsyncGradReducer, ConvBNAct, ConvSpec, EdgeConv, Saved, WeightPacks, _nhwc_empty, act_dtype
from .unet_program import _put

ARCH_KEY = "__upconv_arch__"


def _block_prefix(level: int) -> str:
    """state_dict prefix of the UnetSkipConnectionBlock whose down conv produces e_level (level 1 = outermost)."""
    return "model.model" if level == 1 else "model.model.1" + ".model.3" * (level - 2)


_CAPTURABLE = {"on": None}


def set_capturable(flag: bool) -> None:
    """Capturable training steps (torch.cuda.graph over ``optimize_parameters``): the generator's dropout masks are drawn from
    a call count kept in device memory (same masks as the eager path, call for call).  Also ``GSB_CAPTURABLE=1``."""
    _CAPTURABLE["on"] = bool(flag)


def capturable() -> bool:
    if _CAPTURABLE["on"] is None:
        import os
        return os.environ.get("GSB_CAPTURABLE", "0") == "1"
    return _CAPTURABLE["on"]


class GeneratorProgram:
    def __init__(self, input_nc: int, output_nc: int, ngf: int = 64, use_dropout: bool = False, instance_norm: bool = False,
                 num_downs: int = 8):
        """instance_norm: --norm instance (networks.py:34-35): InstanceNorm2d without parameters, and every convolution
        carries a bias (use_bias, networks.py:568-571).  num_downs: 8 = unet_256, 7 = unet_128 (networks.py:148-151);
        level D = num_downs is the innermost block (layer_index 0), dropout sits on the ngf*8 blocks in between."""
        if not 5 <= num_downs <= 8:
            raise NotImplementedError("genseg_b200: UnetGenerator with num_downs=%d (upconv_arch has 8 rows; the reference "
                                      "builds 7 or 8 levels)" % num_downs)
        self.input_nc, self.output_nc, self.ngf, self.use_dropout = input_nc, output_nc, ngf, use_dropout
        self.inst = inst = bool(instance_norm)
        self.D = D = int(num_downs)
        c = [None] + [ngf * min(2 ** (L - 1), 8) for L in range(1, D + 1)]  # channels of e_1..e_D
        self.c = c
        self.down: List[Optional[ConvBNAct]] = [None]
        for L in range(1, D + 1):
            cin = input_nc if L == 1 else c[L - 1]
            spec = ConvSpec("conv", cin, c[L], 4, 2, 1)
            if L == D:
                self.down.append(ConvBNAct(spec, False, ACT_RELU))
            else:
                self.down.append(ConvBNAct(spec, L != 1, ACT_LRELU, ACT_RELU, instance_norm=inst))
        self.up: List[Optional[ConvBNAct]] = [None] * D
        for L in range(D - 1, 0, -1):
            cin = c[D] if L == D - 1 else 2 * c[L + 1]
            self.up[L] = ConvBNAct(ConvSpec("convT", cin, c[L], 8, 2, 3), True, ACT_RELU,
                                   dropout=use_dropout and 4 <= L < D - 1, instance_norm=inst)
        self.outer = EdgeConv(ConvSpec("convT", 2 * c[1], output_nc, 8, 2, 3), ACT_TANH)
        names: List[str] = ["model.model.0.weight"] + (["model.model.0.bias"] if inst else [])
        for L in range(2, D + 1):
            q = _block_prefix(L)
            names.append(q + ".model.1.weight")
            if inst:
                names.append(q + ".model.1.bias")
            elif L != D:
                names += [q + ".model.2.weight", q + ".model.2.bias"]
        for L in range(D - 1, 0, -1):
            q = _block_prefix(L + 1)
            ui, bi = (3, 4) if L == D - 1 else (5, 6)
            names += [f"{q}.model.{ui}._ops._ops.{i}.op.weight" for i in range(3)]
            if inst:
                names += [f"{q}.model.{ui}._ops._ops.{i}.op.bias" for i in range(3)]
            else:
                names += [f"{q}.model.{bi}.weight", f"{q}.model.{bi}.bias"]
        names += [f"model.model.3._ops._ops.{i}.op.weight" for i in range(3)]
        names += [f"model.model.3._ops._ops.{i}.op.bias" for i in range(3)]
        names.append(ARCH_KEY)
        self.param_names = names
        self._calls = 0
        self._calls_dev = None        # capturable mode: the call count in device memory (int64 scalar)
        self._calls_mirror = None     # the host count the device counter was last known to equal

    @staticmethod
    def _bn(P, B, p: str):
        return (P[p + ".weight"], P[p + ".bias"], B[p + ".running_mean"], B[p + ".running_var"],
                B[p + ".num_batches_tracked"])

    def _up_names(self, L: int):
        q = _block_prefix(L + 1)
        ui, bi = (3, 4) if L == self.D - 1 else (5, 6)
        return [f"{q}.model.{ui}._ops._ops.{i}.op.weight" for i in range(3)], f"{q}.model.{bi}"

    def _up_bias_names(self, L: int):
        q = _block_prefix(L + 1)
        ui = 3 if L == self.D - 1 else 5
        return [f"{q}.model.{ui}._ops._ops.{i}.op.bias" for i in range(3)]

    def _down_bias(self, P, L: int):
        if not self.inst:
            return None
        return P["model.model.0.bias" if L == 1 else _block_prefix(L) + ".model.1.bias"]

    def forward(self, x: torch.Tensor, P, B, training: bool, save: bool, extra=None):
        x = x.contiguous()
        if x.dtype != torch.float32:
            x = x.float()
        n, cin, h, w = x.shape
        D = self.D
        if cin != self.input_nc or h % (1 << D) != 0 or w % (1 << D) != 0:
            raise RuntimeError("UnetGenerator(num_downs=%d): expected input [N,%d,%dk,%dk], got %s"
                               % (D, self.input_nc, 1 << D, 1 << D, tuple(x.shape)))
        dev = x.device
        arch = P[ARCH_KEY]
        if arch.dtype != torch.float32 or tuple(arch.shape) != (8, 3):
            raise RuntimeError("upconv_arch must be a float32 [8,3] tensor")
        arch = arch.contiguous()
        self._calls += 1
        masks = (extra or {}).get("drop_masks")
        # one seed per forward call, the same on every rank (same torch seed, same call count); the masks are indexed by the
        # GLOBAL element index (engine.first_global_sample), so a batch sharded over N ranks draws the one-rank masks
        base_seed = int((extra or {}).get("drop_seed", 0)) or (torch.initial_seed() * 1000003 + self._calls) % (1 << 62) + 1
        if (capturable() and training and masks is None and not int((extra or {}).get("drop_seed", 0))
                and any(self.up[L].dropout for L in range(1, D))):
            # CUDA-graph capturable dropout: the call count lives in device memory and explicit keep masks holding exactly
            # the bits of the seed rule above are drawn from it by a kernel, so every REPLAY of a captured step draws new
            # masks (with the host-side count the capture step's masks would be frozen into the graph)
            if self._calls_dev is None or self._calls_dev.device != dev:
                self._calls_dev = torch.zeros((), device=dev, dtype=torch.int64)
                self._calls_mirror = None
            if self._calls_mirror is None or self._calls != self._calls_mirror + 1:
                self._calls_dev.fill_(self._calls - 1)          # first use, or somebody re-seeded the host count
            ops.counter_add(self._calls_dev, 1)
            self._calls_mirror = self._calls
            seed_k = (torch.initial_seed() * 1000003) % (1 << 62)
            from .engine import first_global_sample
            masks = []
            for L in range(4, D):                               # masks[L - 4] belongs to level L (levels without dropout: None)
                if L < D and self.up[L].dropout:
                    oh, ow = h >> L, w >> L                      # output of up-cell L = spatial size of level L
                    masks.append(ops.dropout_mask((n, oh, ow, self.c[L]), dev, seed_k, self._calls_dev, 7919 * L,
                                                  first_global_sample(n) * oh * ow * self.c[L]))
                else:
                    masks.append(None)
        tape: Dict[str, object] = {"x": x, "arch": arch}
        c = self.c
        sizes = [(h, w)] + [(h >> L, w >> L) for L in range(1, D + 1)]      # sizes[L] = spatial of e_L
        cats = [None] + [_nhwc_empty(n, sizes[L][0], sizes[L][1], 2 * c[L], dev) for L in range(1, D)]
        # ---- weights: the eight merges first, then every tensor-core weight (forward and, when a tape is kept, backward
        # layout) packed by ONE launch instead of one or two per layer (the eager step is host-bound: profiles/README.md 13)
        packs = WeightPacks()
        for L in range(2, D + 1):
            packs.add(f"down{L}", self.down[L].spec, P[_block_prefix(L) + ".model.1.weight"], need_bwd=save)
        merged = {}
        for L in range(D - 1, 0, -1):
            wn, _ = self._up_names(L)
            if self.inst:
                bnm = self._up_bias_names(L)
                merged[L] = ops.merge_upconv(P[wn[0]], P[wn[1]], P[wn[2]], P[bnm[0]], P[bnm[1]], P[bnm[2]], arch[D - 1 - L])
            else:
                merged[L] = (ops.merge_upconv(P[wn[0]], P[wn[1]], P[wn[2]], None, None, None, arch[D - 1 - L])[0], None)
            packs.add(f"up{L}", self.up[L].spec, merged[L][0], need_bwd=save)
        packs.run()
        # ---- encoder
        cur_view, cur_keep, ch, cw = view_nchw(x), x, h, w
        for L in range(1, D + 1):
            q = _block_prefix(L)
            wname = "model.model.0.weight" if L == 1 else q + ".model.1.weight"
            bn = self._bn(P, B, q + ".model.2") if (1 < L < D and not self.inst) else None
            out2 = cats[L][..., :c[L]] if L < D else None
            s, sv = self.down[L].forward(cur_view, cur_keep, n, ch, cw, dev, P[wname], self._down_bias(P, L), bn, training,
                                         out2=out2, save=save, packed=packs.get(f"down{L}") if L > 1 else None)
            tape[f"down{L}"] = sv
            cur_view, cur_keep, (ch, cw) = view_nhwc(s), s, sizes[L]
        # ---- decoder
        cur = cur_keep                                   # u = relu(e_D)
        for L in range(D - 1, 0, -1):
            wn, bnp = self._up_names(L)
            wm, bm_up = merged[L]
            hh, ww = cur.shape[1], cur.shape[2]
            drop = self.up[L].dropout and training
            mask = masks[L - 4] if (drop and masks is not None) else None
            seed = (base_seed + 7919 * L) if (drop and masks is None) else 0
            _, sv = self.up[L].forward(view_nhwc(cur), cur, n, hh, ww, dev, wm, bm_up,
                                       None if self.inst else self._bn(P, B, bnp), training,
                                       out1=cats[L][..., c[L]:], save=save, drop_seed=seed, drop_mask=mask,
                                       packed=packs.get(f"up{L}"))
            tape[f"up{L}"] = sv
            cur = cats[L]
        wo = [f"model.model.3._ops._ops.{i}.op.weight" for i in range(3)]
        bo = [f"model.model.3._ops._ops.{i}.op.bias" for i in range(3)]
        wm, bm = ops.merge_upconv(P[wo[0]], P[wo[1]], P[wo[2]], P[bo[0]], P[bo[1]], P[bo[2]], arch[D - 1])
        out, sv = self.outer.forward(cur, wm, bm, save)
        tape["outer"] = sv
        tape["P"] = P if save else None
        tape["sizes"], tape["n"] = sizes, n
        return out, tape

    def backward(self, tape, dout: torch.Tensor, need: Dict[str, bool], need_dx: bool):
        G: Dict[str, torch.Tensor] = {}
        P = tape["P"]
        arch = tape["arch"]
        dev = dout.device
        c, sizes, n, D = self.c, tape["sizes"], tape["n"], self.D
        dalpha = torch.zeros_like(arch)
        # Data parallelism (parallel.init_data_parallel(overlap_grad_allreduce=True)): the MERGED 8x8 up-cell gradients are
        # all-reduced -- 64 instead of 4*4 + 6*6 + 8*8 = 116 values per (ci, co) pair, asynchronously, under the rest of the
        # backward pass -- and un-merged afterwards: the un-merge (crops, softmax weights, the dot products of d alpha) is
        # linear in dWm, so dW4 / dW6 / dW8 / d alpha come out summed over the ranks.  The down-convolution gradients
        # follow as they are produced; BatchNorm / bias gradients go in one small bucket at the end.
        red = AsyncGradReducer()
        deferred = []

        def unmerge_or_defer(dwm, dbm, wn, bn_, row):
            if not red.active:
                unmerge(dwm, dbm, wn, bn_, row)
                return
            red.start(dwm)
            if dbm is not None:
                red.start(dbm)
            deferred.append((dwm, dbm, wn, bn_, row))

        def unmerge(dwm, dbm, wn, bn_, row):
            ws = [P[k] for k in wn]
            bs = [P[k] for k in bn_] if bn_ else [None, None, None]
            o = ops.unmerge_upconv(dwm, dbm, ws[0], ws[1], ws[2], bs[0], bs[1], bs[2], arch[row], dalpha[row])
            for i in range(3):
                if need.get(wn[i], False):
                    G[wn[i]] = o[i]
                if bn_ and need.get(bn_[i], False):
                    G[bn_[i]] = o[3 + i]

        wo = [f"model.model.3._ops._ops.{i}.op.weight" for i in range(3)]
        bo = [f"model.model.3._ops._ops.{i}.op.bias" for i in range(3)]
        need_arch = need.get(ARCH_KEY, False)
        need_wo = need_arch or any(need.get(k, False) for k in wo + bo)
        dcat, dwm, dbm = self.outer.backward(tape["outer"], dout, True, need_wo, need_wo)
        if need_wo:
            unmerge_or_defer(dwm, dbm, wo, bo, D - 1)
        dskips: List[Optional[torch.Tensor]] = [None] * (D + 1)
        for L in range(1, D):
            wn, bnp = self._up_names(L)
            ubn = self._up_bias_names(L) if self.inst else None
            nw = need_arch or any(need.get(k, False) for k in wn + (ubn or []))
            nb = (not self.inst) and (need.get(bnp + ".weight", False) or need.get(bnp + ".bias", False))
            dskips[L] = dcat[..., :c[L]]
            dnext, dwm, dbm_up, dg, db = self.up[L].backward(tape[f"up{L}"], dcat[..., c[L]:], None, True, nw, nb,
                                                             self.inst and nw)
            if nw:
                unmerge_or_defer(dwm, dbm_up if self.inst else None, wn, ubn, D - 1 - L)
            if not self.inst:
                _put(G, bnp + ".weight", dg)
                _put(G, bnp + ".bias", db)
            dcat = dnext            # gradient of cat_{L+1} (or of u for L == D - 1)
        ds = dcat
        dx = None
        big_done = set()
        for L in range(D, 0, -1):
            q = _block_prefix(L)
            wname = "model.model.0.weight" if L == 1 else q + ".model.1.weight"
            nb = ((need.get(q + ".model.2.weight", False) or need.get(q + ".model.2.bias", False))
                  if (1 < L < D and not self.inst) else False)
            bname = ("model.model.0.bias" if L == 1 else q + ".model.1.bias") if self.inst else None
            want_dx = True if L > 1 else need_dx
            dxo = dxv = None
            if L == 1 and need_dx:
                dxo = torch.empty_like(tape["x"])
                dxv = view_nchw(dxo)
            ds, dw, dbias, dg, db = self.down[L].backward(tape[f"down{L}"], ds, dskips[L] if L < D else None, want_dx,
                                                          need.get(wname, False), nb, bool(bname and need.get(bname, False)),
                                                          dx_out=dxo, dx_view=dxv)
            _put(G, wname, dw)
            if red.active and dw is not None and dw.is_contiguous() and dw.numel() >= (1 << 16):
                red.start(dw)
                big_done.add(wname)
            if bname:
                _put(G, bname, dbias)
            if 1 < L < D and not self.inst:
                _put(G, q + ".model.2.weight", dg)
                _put(G, q + ".model.2.bias", db)
            if L == 1:
                dx = ds
        if red.active:
            red.finish([g for k, g in G.items() if k not in big_done])     # BatchNorm / bias / first-layer gradients
            for dwm, dbm, wn, bn_, row in deferred:                          # summed merged gradients -> dW4/6/8, d alpha
                unmerge(dwm, dbm, wn, bn_, row)
            G["__reduced__"] = True
        if need_arch:
            G[ARCH_KEY] = dalpha
        return dx, G


class DiscriminatorProgram:
    """NLayerDiscriminator(input_nc, ndf, n_layers, norm): 'basic' = 3 layers (70x70 patches), 'n_layers' takes
    --n_layers_D (networks.py:197-200).  Sequential indices: conv 0, then conv/norm/LeakyReLU triples at 2, 5, ...,
    2 + 3 (n_layers - 1) (the last with stride 1), and the 1-channel head at 2 + 3 n_layers."""

    def __init__(self, input_nc: int, ndf: int = 64, n_layers: int = 3, instance_norm: bool = False):
        if n_layers < 1:
            raise ValueError("NLayerDiscriminator needs n_layers >= 1")
        self.input_nc = input_nc
        self.inst = inst = bool(instance_norm)
        ch = [ndf * min(2 ** i, 8) for i in range(n_layers + 1)]
        self.l0 = ConvBNAct(ConvSpec("conv", input_nc, ndf, 4, 2, 1), False, ACT_LRELU)
        self.mid = [(2 + 3 * (i - 1), ConvBNAct(ConvSpec("conv", ch[i - 1], ch[i], 4, 2 if i < n_layers else 1, 1), True, ACT_LRELU,
                                                instance_norm=inst)) for i in range(1, n_layers + 1)]
        self.head = 2 + 3 * n_layers
        self.last = EdgeConv(ConvSpec("conv", ch[n_layers], 1, 4, 1, 1), ACT_NONE)
        names = ["model.0.weight", "model.0.bias"]
        for idx, _ in self.mid:
            if inst:    # InstanceNorm2d(affine=False): no norm parameters, the convolution has a bias
                names += [f"model.{idx}.weight", f"model.{idx}.bias"]
            else:
                names += [f"model.{idx}.weight", f"model.{idx + 1}.weight", f"model.{idx + 1}.bias"]
        names += [f"model.{self.head}.weight", f"model.{self.head}.bias"]
        self.param_names = names

    def forward(self, x: torch.Tensor, P, B, training: bool, save: bool, extra=None):
        x = x.contiguous()
        if x.dtype != torch.float32:
            x = x.float()
        n, cin, h, w = x.shape
        if cin != self.input_nc:
            raise RuntimeError("NLayerDiscriminator: expected %d input channels, got %d" % (self.input_nc, cin))
        dev = x.device
        tape: Dict[str, object] = {"x": x}
        a, sv = self.l0.forward(view_nchw(x), x, n, h, w, dev, P["model.0.weight"], P["model.0.bias"], None, training,
                                save=save)
        tape["l0"] = sv
        packs = WeightPacks()
        for idx, layer in self.mid:
            packs.add(f"l{idx}", layer.spec, P[f"model.{idx}.weight"], need_bwd=save)
        packs.run()
        for idx, layer in self.mid:
            p = f"model.{idx + 1}"
            bn = None if self.inst else (P[p + ".weight"], P[p + ".bias"], B[p + ".running_mean"], B[p + ".running_var"],
                                         B[p + ".num_batches_tracked"])
            a, sv = layer.forward(view_nhwc(a), a, n, a.shape[1], a.shape[2], dev, P[f"model.{idx}.weight"],
                                  P[f"model.{idx}.bias"] if self.inst else None, bn, training, save=save,
                                  packed=packs.get(f"l{idx}"))
            tape[f"l{idx}"] = sv
        out, sv = self.last.forward(a, P[f"model.{self.head}.weight"], P[f"model.{self.head}.bias"], save)
        tape["last"] = sv
        return out, tape

    def backward(self, tape, dout, need: Dict[str, bool], need_dx: bool):
        G: Dict[str, torch.Tensor] = {}
        hw_, hb_ = f"model.{self.head}.weight", f"model.{self.head}.bias"
        d, dw, db = self.last.backward(tape["last"], dout, True, need.get(hw_, False), need.get(hb_, False))
        _put(G, hw_, dw)
        _put(G, hb_, db)
        for idx, layer in reversed(self.mid):
            p = f"model.{idx + 1}"
            nb = (not self.inst) and (need.get(p + ".weight", False) or need.get(p + ".bias", False))
            d, dw, dcb, dg, dbb = layer.backward(tape[f"l{idx}"], d, None, True, need.get(f"model.{idx}.weight", False), nb,
                                                 self.inst and need.get(f"model.{idx}.bias", False))
            _put(G, f"model.{idx}.weight", dw)
            if self.inst:
                _put(G, f"model.{idx}.bias", dcb)
            else:
                _put(G, p + ".weight", dg)
                _put(G, p + ".bias", dbb)
        dx = dxv = None
        if need_dx:
            dx = torch.empty_like(tape["x"])
            dxv = view_nchw(dx)
        _, dw, db, _, _ = self.l0.backward(tape["l0"], d, None, need_dx, need.get("model.0.weight", False), False,
                                           need.get("model.0.bias", False), dx_out=dx, dx_view=dxv)
        _put(G, "model.0.weight", dw)
        _put(G, "model.0.bias", db)
        return dx, G
