"""Python wrappers over the C ABI (one function per entry point family).

Tensors are torch CUDA tensors used as typed device buffers; activations inside the engine are
NHWC ([N,H,W,C], bf16 in production / fp32 in validation mode), possibly channel slices of
concat buffers.  Everything is enqueued on torch's current CUDA stream.
"""
from __future__ import annotations

import ctypes as C
from typing import Optional, Tuple

import torch

from . import _lib as L
from ._lib import (ACT_LRELU, ACT_NONE, ACT_RELU, ACT_TANH, MODE_CONV, MODE_CONVT, GsbConvGeom, GsbView, check,
                   load, ptr, stream_ptr, view_nchw, view_nhwc)

__all__ = ["ACT_NONE", "ACT_RELU", "ACT_LRELU", "ACT_TANH", "MODE_CONV", "MODE_CONVT"]


def _s(t: torch.Tensor) -> int:
    return stream_ptr(t.device)


_ARRAY_TYPES = {}


def _atype(ctype, n: int):
    """ctypes array types are created on demand and only weakly cached: keep them, or every call leaves a few cyclic
    type objects behind for the garbage collector."""
    key = (ctype, n)
    t = _ARRAY_TYPES.get(key)
    if t is None:
        t = _ARRAY_TYPES[key] = ctype * n
    return t


def _ref(v: Optional[GsbView]):
    return C.byref(v) if v is not None else None


# Per-(device, stream) scratch for kernels whose partial results are consumed INSIDE the same C call (per-CTA partial rows of
# the deterministic two-stage reductions, split-K slices of the small-tile convolution).  Calls on one stream are ordered,
# so the buffer is reused; a buffer that had to grow is kept alive (a captured CUDA graph may still hold its address).
_SCRATCH = {}
_SCRATCH_RETIRED = []


def _scratch(device, nbytes: int) -> torch.Tensor:
    idx = device.index if isinstance(device, torch.device) else int(device)
    if idx is None:
        idx = torch.cuda.current_device()
    key = (idx, stream_ptr(idx))
    buf = _SCRATCH.get(key)
    if buf is None or buf.numel() < nbytes:
        if buf is not None:
            _SCRATCH_RETIRED.append(buf)
        buf = torch.empty(max(int(nbytes), 1 << 20), device=torch.device("cuda", idx), dtype=torch.uint8)
        _SCRATCH[key] = buf
    return buf


# ---- layout ---------------------------------------------------------------------------------

def nchw_to_nhwc(x: torch.Tensor, dtype: torch.dtype) -> torch.Tensor:
    n, c, h, w = x.shape
    out = torch.empty((n, h, w, c), device=x.device, dtype=dtype)
    check(load().gsb_copy_view(_ref(view_nchw(x)), _ref(view_nhwc(out)), _s(x)), "copy_view")
    return out


def nhwc_to_nchw(x: torch.Tensor, dtype: torch.dtype = torch.float32) -> torch.Tensor:
    n, h, w, c = x.shape
    out = torch.empty((n, c, h, w), device=x.device, dtype=dtype)
    check(load().gsb_copy_view(_ref(view_nhwc(x)), _ref(view_nchw(out)), _s(x)), "copy_view")
    return out


def copy_view(src: GsbView, dst: GsbView, stream: int) -> None:
    check(load().gsb_copy_view(_ref(src), _ref(dst), stream), "copy_view")


def cast_flat(src: torch.Tensor, dst: torch.Tensor) -> None:
    """dst <- cast(src) for contiguous fp32 / bf16 buffers of equal length (gradient buckets)."""
    assert src.is_contiguous() and dst.is_contiguous() and src.numel() == dst.numel()
    check(load().gsb_cast_flat(ptr(src), L.dtype_code(src.dtype), ptr(dst), L.dtype_code(dst.dtype), src.numel(), _s(src)),
          "cast_flat")


def pack_weight(w: torch.Tensor, rows_from_d0: bool, dtype: torch.dtype) -> torch.Tensor:
    """fp32 [d0,d1,kh,kw] -> [kh*kw, rows, cols] in `dtype` (cols contiguous)."""
    assert w.dtype == torch.float32 and w.is_contiguous()
    d0, d1, kh, kw = w.shape
    rows, cols = (d0, d1) if rows_from_d0 else (d1, d0)
    out = torch.empty((kh * kw, rows, cols), device=w.device, dtype=dtype)
    check(load().gsb_pack_weight(ptr(w), d0, d1, kh, kw, int(rows_from_d0), ptr(out), L.dtype_code(dtype), _s(w)),
          "pack_weight")
    return out


def unpack_wgrad(dwp: torch.Tensor, shape: Tuple[int, int, int, int], rows_from_d0: bool,
                 out: Optional[torch.Tensor] = None, nslices: int = 1) -> torch.Tensor:
    """dwp: fp32 [nslices, kh*kw, rows, cols] partial-sum slices (or one slice), added in slice order while un-packing."""
    d0, d1, kh, kw = shape
    acc = out is not None
    if out is None:
        out = torch.empty(shape, device=dwp.device, dtype=torch.float32)
    check(load().gsb_unpack_wgrad(ptr(dwp), int(nslices), d0, d1, kh, kw, int(rows_from_d0), ptr(out), int(acc), _s(dwp)),
          "unpack_wgrad")
    return out


def sum_slices(ws: torch.Tensor, nslices: int, n: int, dst: torch.Tensor, accumulate: bool = False) -> None:
    check(load().gsb_sum_slices(ptr(ws), int(nslices), int(n), ptr(dst), int(accumulate), _s(ws)), "sum_slices")


def _ptr_array(ts):
    return _atype(C.c_void_p, len(ts))(*[t.data_ptr() for t in ts])


def _int_array(vs):
    return _atype(C.c_int, len(vs))(*[int(v) for v in vs])


def pack_weight_multi(jobs, dtype: torch.dtype) -> None:
    """jobs: [(w fp32 [d0,d1,kh,kw], rows_from_d0, out [kh*kw, rows, cols] in dtype)] -- one launch for all of them."""
    if not jobs:
        return
    ws = [j[0] for j in jobs]
    for w, _, out in jobs:
        assert w.dtype == torch.float32 and w.is_contiguous() and out.dtype == dtype and out.is_contiguous()
        L.require_cuda(w)
    check(load().gsb_pack_weight_multi(len(jobs), _ptr_array(ws), _ptr_array([j[2] for j in jobs]),
                                       _int_array([w.shape[0] for w in ws]), _int_array([w.shape[1] for w in ws]),
                                       _int_array([w.shape[2] * w.shape[3] for w in ws]),
                                       _int_array([int(j[1]) for j in jobs]), L.dtype_code(dtype), _s(ws[0])),
          "pack_weight_multi")


def unpack_wgrad_multi(jobs) -> None:
    """jobs: [(dwp fp32 [nslices, kh*kw, rows, cols], rows_from_d0, out fp32 [d0,d1,kh,kw], nslices)] -- one launch; the
    partial-sum slices of every job are added in slice order while un-packing."""
    if not jobs:
        return
    outs = [j[2] for j in jobs]
    for dwp, _, out, _ns in jobs:
        assert dwp.dtype == torch.float32 and out.dtype == torch.float32 and dwp.is_contiguous() and out.is_contiguous()
        L.require_cuda(out)
    check(load().gsb_unpack_wgrad_multi(len(jobs), _ptr_array([j[0] for j in jobs]), _int_array([j[3] for j in jobs]),
                                        _ptr_array(outs),
                                        _int_array([t.shape[0] for t in outs]), _int_array([t.shape[1] for t in outs]),
                                        _int_array([t.shape[2] * t.shape[3] for t in outs]),
                                        _int_array([int(j[1]) for j in jobs]), 0, _s(outs[0])), "unpack_wgrad_multi")


def merge_upconv(w4, w6, w8, b4, b6, b8, alpha_row: torch.Tensor):
    cin, cout = w8.shape[0], w8.shape[1]
    wm = torch.empty_like(w8)
    bm = torch.empty_like(b8) if b8 is not None else None
    check(load().gsb_merge_upconv(ptr(w4), ptr(w6), ptr(w8), ptr(b4), ptr(b6), ptr(b8), ptr(alpha_row), cin, cout,
                                  ptr(wm), ptr(bm), _s(w8)), "merge_upconv")
    return wm, bm


def unmerge_upconv(dwm, dbm, w4, w6, w8, b4, b6, b8, alpha_row, dalpha_row):
    cin, cout = w8.shape[0], w8.shape[1]
    dw4, dw6, dw8 = torch.empty_like(w4), torch.empty_like(w6), torch.empty_like(w8)
    db4 = db6 = db8 = None
    if dbm is not None:
        db4, db6, db8 = torch.empty_like(b4), torch.empty_like(b6), torch.empty_like(b8)
    ws = _scratch(w8.device, 4096 * 3 * 8)      # one row of 3 fp64 partial dot products per CTA
    check(load().gsb_unmerge_upconv(ptr(dwm), ptr(dbm), ptr(w4), ptr(w6), ptr(w8), ptr(b4), ptr(b6), ptr(b8),
                                    ptr(alpha_row), cin, cout, ptr(dw4), ptr(dw6), ptr(dw8), ptr(db4), ptr(db6),
                                    ptr(db8), ptr(dalpha_row), ptr(ws), 4096 * 3 * 8, _s(w8)), "unmerge_upconv")
    return dw4, dw6, dw8, db4, db6, db8


# ---- convolution ----------------------------------------------------------------------------

def conv_gather(inp: GsbView, wp: torch.Tensor, bias: Optional[torch.Tensor], out: GsbView, mode: int, k: int,
                stride: int, pad: int, act: int = ACT_NONE, accumulate: bool = False, use_tc: int = 2,
                want_stats: bool = False, stream: Optional[int] = None):
    """Returns (stats_partial, rows) when want_stats and the tcgen05 path ran, else (None, 0)."""
    lib = load()
    g = GsbConvGeom(mode, k, k, stride, pad)
    stats = None
    rows = C.c_int(0)
    if want_stats and use_tc != 0 and wp.dtype == torch.bfloat16:
        nrows = lib.gsb_conv_stats_rows(_ref(out), C.byref(g))
        stats = torch.empty((nrows, 2, out.c), device=wp.device, dtype=torch.float32)
    scratch, sbytes = None, 0
    if use_tc != 0 and wp.dtype == torch.bfloat16 and out.n * out.h * out.w <= 9472 and stream is None:
        # launches with a handful of tiles split their K loop over CTAs into fp32 slices of this buffer (<= 9.7 MB needed)
        scratch, sbytes = _scratch(wp.device, _SPLITK_BYTES), _SPLITK_BYTES
    elif (use_tc != 0 and wp.dtype == torch.bfloat16 and mode == L.MODE_CONVT and k == 8 and stride == 2 and out.c == 1
          and inp.c in (64, 128) and stream is None):
        # the generator's CW -> 1 transposed 8x8 cell on the tensor cores (thin_tc.cu): Z[pixel][64 taps] fp32 between its
        # GEMM and its col2im pass
        sbytes = inp.n * inp.h * inp.w * 64 * 4
        scratch = _scratch(wp.device, sbytes)
    check(lib.gsb_conv_gather(_ref(inp), ptr(wp), L.dtype_code(wp.dtype), ptr(bias), _ref(out), C.byref(g), act,
                              int(accumulate), use_tc, ptr(stats), C.byref(rows), ptr(scratch), sbytes,
                              stream if stream is not None else _s(wp)), "conv_gather")
    if stats is not None and rows.value > 0:
        return stats, rows.value
    return None, 0


_SPLITK_BYTES = 16 << 20


def stem_conv_fwd_affine(x: torch.Tensor, weight: torch.Tensor, scale: torch.Tensor, shift: torch.Tensor, y: torch.Tensor, k: int,
                         stride: int, pad: int, act: int) -> bool:
    """y = act(conv(x) * scale + shift) on the fp32 NCHW network input in one pass (inference); False = not applicable."""
    n, cin, h, w = x.shape
    rc = load().gsb_stem_conv_fwd_affine(ptr(x), n, cin, h, w, ptr(weight), ptr(scale), ptr(shift), _ref(view_nhwc(y)), k, stride, pad,
                                         act, _s(x))
    if rc == 1:
        return False
    check(rc, "stem_conv_fwd_affine")
    return True


def conv_gather_affine(inp: GsbView, wp: torch.Tensor, scale: torch.Tensor, shift: torch.Tensor, out: GsbView, mode: int, k: int,
                       stride: int, pad: int, act: int) -> bool:
    """out = act(conv(inp) * scale + shift) in one pass (inference: BatchNorm folded into the convolution's epilogue).
    False when the shape has no such kernel (the caller then runs conv_gather + bn_act)."""
    g = GsbConvGeom(mode, k, k, stride, pad)
    rc = load().gsb_conv_gather_affine(_ref(inp), ptr(wp), L.dtype_code(wp.dtype), ptr(scale), ptr(shift), _ref(out), C.byref(g), act,
                                       _s(wp))
    if rc == 1:
        return False
    check(rc, "conv_gather_affine")
    return True


def conv_wgrad_plan(inp: GsbView, dout: GsbView, mode: int, k: int, stride: int, pad: int, use_tc: int = 2) -> int:
    """Number of partial-sum slices gsb_conv_wgrad will write for this geometry (host-side, no launch)."""
    g = GsbConvGeom(mode, k, k, stride, pad)
    n = load().gsb_conv_wgrad_plan(_ref(inp), _ref(dout), C.byref(g), use_tc)
    if n < 1:
        check(-1, "conv_wgrad_plan")
    return int(n)


def conv_wgrad_slices(inp: GsbView, dout: GsbView, ws: torch.Tensor, max_slices: int, mode: int, k: int, stride: int,
                      pad: int, use_tc: int = 2) -> int:
    """Weight gradient as partial-sum slices ws[slice][k*k][cout][cin] (plain stores, no atomics); returns the slices used."""
    g = GsbConvGeom(mode, k, k, stride, pad)
    used = C.c_int(0)
    check(load().gsb_conv_wgrad(_ref(inp), _ref(dout), ptr(ws), int(max_slices), C.byref(used), C.byref(g), use_tc, _s(ws)),
          "conv_wgrad")
    return int(used.value)


def conv_wgrad_into(inp: GsbView, dout: GsbView, dwp: torch.Tensor, mode: int, k: int, stride: int, pad: int,
                    use_tc: int = 2) -> None:
    """dwp[k*k][cout][cin] (fp32) += weight gradient: slices into a temporary workspace, added to dwp in slice order."""
    n = conv_wgrad_plan(inp, dout, mode, k, stride, pad, use_tc)
    ws = torch.empty((n,) + tuple(dwp.shape), device=dwp.device, dtype=torch.float32)
    used = conv_wgrad_slices(inp, dout, ws, n, mode, k, stride, pad, use_tc)
    sum_slices(ws, used, dwp.numel(), dwp, accumulate=True)


def stem_conv_fwd(x: torch.Tensor, weight: torch.Tensor, bias, y: torch.Tensor, k: int, stride: int, pad: int,
                  act: int = ACT_NONE, want_stats: bool = False):
    """Returns (stats_partial, rows) when want_stats and the fused-statistics kernel ran, else (None, 0)."""
    n, cin, h, w = x.shape
    lib = load()
    stats = None
    rows = C.c_int(0)
    if want_stats:
        stats = torch.empty((lib.gsb_stem_stats_rows(), 2, y.shape[3]), device=x.device, dtype=torch.float32)
    check(lib.gsb_stem_conv_fwd(ptr(x), n, cin, h, w, ptr(weight), ptr(bias), _ref(view_nhwc(y)), k, stride, pad, act,
                                ptr(stats), C.byref(rows), _s(x)), "stem_conv_fwd")
    if stats is not None and rows.value > 0:
        return stats, rows.value
    return None, 0


def stem_conv_wgrad(x: torch.Tensor, dy: torch.Tensor, k: int, stride: int, pad: int) -> torch.Tensor:
    n, cin, h, w = x.shape
    dw = torch.empty((dy.shape[3], cin, k, k), device=x.device, dtype=torch.float32)
    nbytes = load().gsb_edge_rows() * dw.numel() * 4          # one partial row per CTA column, added in row order
    ws = _scratch(x.device, nbytes)
    check(load().gsb_stem_conv_wgrad(ptr(x), n, cin, h, w, _ref(view_nhwc(dy)), k, stride, pad, ptr(dw), ptr(ws), nbytes, _s(x)),
          "stem_conv_wgrad")
    return dw


def head_conv_fwd(x: torch.Tensor, weight: torch.Tensor, bias) -> torch.Tensor:
    n, h, w, _ = x.shape
    c = weight.shape[0]
    out = torch.empty((n, c, h, w), device=x.device, dtype=torch.float32)
    check(load().gsb_head_conv_fwd(_ref(view_nhwc(x)), ptr(weight), ptr(bias), c, ptr(out), _s(x)), "head_conv_fwd")
    return out


def head_conv_bwd(x: torch.Tensor, weight: torch.Tensor, dlogits: torch.Tensor, need_dx: bool, need_dw: bool,
                  need_db: bool):
    c = weight.shape[0]
    dx = torch.empty_like(x) if need_dx else None
    dw = torch.empty_like(weight) if need_dw else None
    db = torch.empty(c, device=x.device, dtype=torch.float32) if need_db else None
    ws, nbytes = None, 0
    if need_dw or need_db:
        nbytes = load().gsb_edge_rows() * (c * x.shape[3] + 8) * 4
        ws = _scratch(x.device, nbytes)
    check(load().gsb_head_conv_bwd(_ref(view_nhwc(x)), ptr(weight), c, ptr(dlogits), _ref(view_nhwc(dx)) if need_dx else None,
                                   ptr(dw), ptr(db), ptr(ws), nbytes, _s(x)), "head_conv_bwd")
    return dx, dw, db


def bias_grad(dout: GsbView, db: torch.Tensor) -> None:
    """db[c] = sum over pixels of dout[., c] (stored)."""
    nbytes = load().gsb_bias_grad_workspace_bytes(_ref(dout))
    ws = _scratch(db.device, nbytes)
    check(load().gsb_bias_grad(_ref(dout), ptr(db), ptr(ws), nbytes, _s(db)), "bias_grad")


# ---- batch norm -----------------------------------------------------------------------------

def _bn_ws(v: GsbView, device) -> torch.Tensor:
    nbytes = load().gsb_bn_workspace_bytes(_ref(v))
    return torch.empty(nbytes // 4, device=device, dtype=torch.float32)


def bn_reduce(x: torch.Tensor, per_image: bool = False) -> torch.Tensor:
    """[sum x, sum x^2] per channel (fp64 [2c]); per_image: per (image, channel), fp64 [2 * n * c] (InstanceNorm)."""
    v = view_nhwc(x)
    ws = _bn_ws(v, x.device)
    sums = torch.empty(2 * v.c * (v.n if per_image else 1), device=x.device, dtype=torch.float64)
    check(load().gsb_bn_reduce(_ref(v), int(per_image), ptr(sums), ptr(ws), ws.numel() * 4, _s(x)), "bn_reduce")
    return sums


def bn_reduce_partials(partial: torch.Tensor, rows: int, c: int) -> torch.Tensor:
    sums = torch.empty(2 * c, device=partial.device, dtype=torch.float64)
    check(load().gsb_bn_reduce_partials(ptr(partial), rows, c, ptr(sums), _s(partial)), "bn_reduce_partials")
    return sums


def bn_finalize(sums: Optional[torch.Tensor], count: float, gamma, beta, running_mean, running_var, nbt,
                momentum: float, eps: float, c: int, device, want_saved: bool = True, centered: bool = False):
    """centered: shift = beta and the consumers get center = mean, i.e. (x - mean) * scale + beta (fp32 validation mode);
    otherwise the folded form x * scale + (beta - mean * scale)."""
    scale = torch.empty(c, device=device, dtype=torch.float32)
    shift = torch.empty(c, device=device, dtype=torch.float32)
    mean = torch.empty(c, device=device, dtype=torch.float32) if want_saved else None
    invstd = torch.empty(c, device=device, dtype=torch.float32) if want_saved else None
    check(load().gsb_bn_finalize(ptr(sums), float(count), ptr(gamma), ptr(beta), ptr(running_mean), ptr(running_var),
                                 ptr(nbt), float(momentum), float(eps), c, ptr(scale), ptr(shift), ptr(mean),
                                 ptr(invstd), int(centered), stream_ptr(device)), "bn_finalize")
    return scale, shift, mean, invstd


def bn_act(x: torch.Tensor, scale, shift, act1: int, out1: torch.Tensor, act2: int = ACT_NONE,
           out2: Optional[torch.Tensor] = None, drop_seed: int = 0, drop_mask: Optional[torch.Tensor] = None,
           center: Optional[torch.Tensor] = None, drop_offset: int = 0, per_image: bool = False) -> None:
    """z = (x - center) * scale + shift (center None = 0); drop_offset: global element index of this tensor's first element
    (Philox dropout masks of a sharded batch equal those of the one-rank run)."""
    v2 = view_nhwc(out2) if out2 is not None else None
    check(load().gsb_bn_act(_ref(view_nhwc(x)), ptr(scale), ptr(shift), ptr(center), act1, _ref(view_nhwc(out1)), act2, _ref(v2),
                            int(drop_seed), int(drop_offset), ptr(drop_mask), int(per_image), _s(x)), "bn_act")


def bn_act_pool(x: torch.Tensor, scale, shift, out1: torch.Tensor, pooled: torch.Tensor) -> None:
    """out1 = relu(x * scale + shift), pooled = MaxPool2d(2)(out1) in one pass (even H and W)."""
    check(load().gsb_bn_act_pool(_ref(view_nhwc(x)), ptr(scale), ptr(shift), _ref(view_nhwc(out1)), _ref(view_nhwc(pooled)), _s(x)),
          "bn_act_pool")


def bn_bwd_reduce(x, dA1, dA2, scale, shift, mean, invstd, act1, act2, drop_seed=0, drop_mask=None, center=None,
                  drop_offset=0) -> torch.Tensor:
    v = view_nhwc(x)
    ws = _bn_ws(v, x.device)
    sums = torch.empty(2 * v.c, device=x.device, dtype=torch.float64)
    v2 = view_nhwc(dA2) if dA2 is not None else None
    check(load().gsb_bn_bwd_reduce(_ref(v), _ref(view_nhwc(dA1)), _ref(v2), ptr(scale), ptr(shift), ptr(center), ptr(mean),
                                   ptr(invstd), act1, act2, int(drop_seed), int(drop_offset), ptr(drop_mask), ptr(sums),
                                   ptr(ws), ws.numel() * 4, _s(x)), "bn_bwd_reduce")
    return sums


def bn_bwd_apply(x, dA1, dA2, scale, shift, mean, invstd, act1, act2, sums, count, dx, dgamma=None, dbeta=None,
                 drop_seed=0, drop_mask=None, center=None, drop_offset=0, per_image=False) -> None:
    v2 = view_nhwc(dA2) if dA2 is not None else None
    check(load().gsb_bn_bwd_apply(_ref(view_nhwc(x)), _ref(view_nhwc(dA1)), _ref(v2), ptr(scale), ptr(shift), ptr(center),
                                  ptr(mean), ptr(invstd), act1, act2, int(drop_seed), int(drop_offset), ptr(drop_mask), ptr(sums),
                                  float(count), _ref(view_nhwc(dx)), ptr(dgamma), ptr(dbeta), int(per_image), _s(x)), "bn_bwd_apply")


def set_sm_reserve(n: int) -> int:
    """SMs set aside for a concurrently running collective (persistent kernels size their grids for the rest)."""
    return int(load().gsb_set_sm_reserve(int(n)))


def _peer_args(peer):
    """peer = None (single GPU) or {"regions": [...], "rank": r, "world": n} from parallel.py."""
    if peer is None:
        return None, 0, 1
    world = int(peer["world"])
    arr = _atype(C.c_void_p, world)(*[C.c_void_p(r) for r in peer["regions"]])
    return arr, int(peer["rank"]), world


def bn_stats_finalize(partial: torch.Tensor, rows: int, c: int, count: float, gamma, beta, running_mean, running_var, nbt,
                      momentum: float, eps: float, peer=None, centered: bool = False):
    """Statistic rows -> (cross-rank sum over the peer regions) -> scale, shift, mean, invstd + running statistics:
    one launch.  ``count`` is the GLOBAL element count per channel."""
    dev = partial.device
    scale = torch.empty(c, device=dev, dtype=torch.float32)
    shift = torch.empty(c, device=dev, dtype=torch.float32)
    mean = torch.empty(c, device=dev, dtype=torch.float32)
    invstd = torch.empty(c, device=dev, dtype=torch.float32)
    arr, rank, world = _peer_args(peer)
    check(load().gsb_bn_stats_finalize(ptr(partial), int(rows), c, float(count), arr, rank, world, ptr(gamma), ptr(beta),
                                       ptr(running_mean), ptr(running_var), ptr(nbt), float(momentum), float(eps), ptr(scale),
                                       ptr(shift), ptr(mean), ptr(invstd), int(centered), _s(partial)), "bn_stats_finalize")
    return scale, shift, mean, invstd


def bn_bwd_stats(x, dA1, dA2, scale, shift, mean, invstd, act1, act2, drop_seed=0, drop_mask=None, want_param_grads=True,
                 peer=None, center=None, drop_offset=0, per_image=False, pool_grad=None):
    """[sum dy, sum dy*xhat] in two launches: per-block rows, then totals + dgamma / dbeta (local) + cross-rank sum.
    -> (global sums fp64 [2c], dgamma, dbeta).  pool_grad (BatchNorm + ReLU layers in front of a MaxPool2d(2), bf16, even
    sizes): the pooled gradient is first routed into dA1 IN PLACE by the same pass (the max-pool backward)."""
    v = view_nhwc(x)
    ws = _bn_ws(v, x.device)
    rows = C.c_int(0)
    v2 = view_nhwc(dA2) if dA2 is not None else None
    if pool_grad is not None:
        check(load().gsb_bn_bwd_partial_pool(_ref(v), _ref(view_nhwc(dA1)), _ref(view_nhwc(pool_grad)), ptr(scale), ptr(shift),
                                             ptr(mean), ptr(invstd), ptr(ws), ws.numel() * 4, C.byref(rows), _s(x)),
              "bn_bwd_partial_pool")
    else:
        check(load().gsb_bn_bwd_partial(_ref(v), _ref(view_nhwc(dA1)), _ref(v2), ptr(scale), ptr(shift), ptr(center), ptr(mean),
                                        ptr(invstd), act1, act2, int(drop_seed), int(drop_offset), ptr(drop_mask), ptr(ws),
                                        ws.numel() * 4, C.byref(rows), int(per_image), _s(x)), "bn_bwd_partial")
    cc = v.c * (v.n if per_image else 1)          # per_image: one "channel" per (image, channel)
    sums = torch.empty(2 * cc, device=x.device, dtype=torch.float64)
    dgamma = torch.empty(cc, device=x.device, dtype=torch.float32) if want_param_grads else None
    dbeta = torch.empty(cc, device=x.device, dtype=torch.float32) if want_param_grads else None
    arr, rank, world = _peer_args(peer)
    check(load().gsb_bn_bwd_stats(ptr(ws), rows.value, cc, arr, rank, world, ptr(dgamma), ptr(dbeta), 0, ptr(sums), _s(x)),
          "bn_bwd_stats")
    return sums, dgamma, dbeta


def bn_param_grad(sums: torch.Tensor, c: int):
    dgamma = torch.empty(c, device=sums.device, dtype=torch.float32)
    dbeta = torch.empty(c, device=sums.device, dtype=torch.float32)
    check(load().gsb_bn_param_grad(ptr(sums), c, ptr(dgamma), ptr(dbeta), 0, _s(sums)), "bn_param_grad")
    return dgamma, dbeta


# ---- pooling / resampling -------------------------------------------------------------------

def maxpool2_fwd(x: torch.Tensor) -> torch.Tensor:
    n, h, w, c = x.shape
    y = torch.empty((n, h // 2, w // 2, c), device=x.device, dtype=x.dtype)
    check(load().gsb_maxpool2_fwd(_ref(view_nhwc(x)), _ref(view_nhwc(y)), _s(x)), "maxpool2_fwd")
    return y


def maxpool2_bwd(x: torch.Tensor, dy: torch.Tensor, dx: torch.Tensor, accumulate: bool) -> None:
    check(load().gsb_maxpool2_bwd(_ref(view_nhwc(x)), _ref(view_nhwc(dy)), _ref(view_nhwc(dx)), int(accumulate),
                                  _s(x)), "maxpool2_bwd")


def upsample2_fwd(x: torch.Tensor, out: torch.Tensor, pad_top: int, pad_left: int) -> None:
    check(load().gsb_upsample2_fwd(_ref(view_nhwc(x)), _ref(view_nhwc(out)), pad_top, pad_left, _s(x)),
          "upsample2_fwd")


def upsample2_bwd(dout: torch.Tensor, dx: torch.Tensor, pad_top: int, pad_left: int) -> None:
    check(load().gsb_upsample2_bwd(_ref(view_nhwc(dout)), _ref(view_nhwc(dx)), pad_top, pad_left, _s(dx)),
          "upsample2_bwd")


# ---- losses ---------------------------------------------------------------------------------

def seg_loss_reduce(logits: torch.Tensor, labels: torch.Tensor) -> torch.Tensor:
    n, c, h, w = logits.shape
    lib = load()
    nbytes = lib.gsb_seg_loss_workspace_bytes(n, c, h, w)
    ws = torch.empty(nbytes // 8, device=logits.device, dtype=torch.float64)
    sums = torch.empty(4, device=logits.device, dtype=torch.float64)
    check(lib.gsb_seg_loss_reduce(ptr(logits), ptr(labels), n, c, h, w, ptr(sums), ptr(ws), nbytes, _s(logits)),
          "seg_loss_reduce")
    return sums


def seg_loss_finalize(sums: torch.Tensor, count: float) -> torch.Tensor:
    out = torch.empty(3, device=sums.device, dtype=torch.float32)
    check(load().gsb_seg_loss_finalize(ptr(sums), float(count), ptr(out), _s(sums)), "seg_loss_finalize")
    return out


def seg_loss_bwd(logits, labels, sums, count: float, gscale: torch.Tensor) -> torch.Tensor:
    n, c, h, w = logits.shape
    d = torch.empty_like(logits)
    check(load().gsb_seg_loss_bwd(ptr(logits), ptr(labels), n, c, h, w, ptr(sums), float(count), ptr(gscale), ptr(d),
                                  _s(logits)), "seg_loss_bwd")
    return d


def eval_counts(logits: torch.Tensor, labels: torch.Tensor, want_mask: bool = True):
    n, c, h, w = logits.shape
    k = max(c - 1, 1)
    counts = torch.empty((n, k, 3), device=logits.device, dtype=torch.int64)
    mask = torch.empty((n, h, w), device=logits.device, dtype=torch.uint8) if want_mask else None
    check(load().gsb_eval_counts(ptr(logits), ptr(labels), n, c, h, w, ptr(mask), ptr(counts), _s(logits)),
          "eval_counts")
    return counts, mask


def dice_sums(a: torch.Tensor, b: torch.Tensor, slices: int, inner: int) -> torch.Tensor:
    sums = torch.empty((slices, 3), device=a.device, dtype=torch.float64)
    rows = max(1, min(1024, (inner + 4095) // 4096, (8 << 20) // (slices * 24)))
    nbytes = rows * slices * 24
    ws = _scratch(a.device, nbytes)
    check(load().gsb_dice_sums(ptr(a), ptr(b), slices, inner, ptr(sums), ptr(ws), nbytes, _s(a)), "dice_sums")
    return sums


_LOSS_WS_BYTES = 2048 * 8      # one fp64 partial per CTA (grids are capped at 148 * 8 CTAs)


def gan_loss_fwd(pred: torch.Tensor, mode: int, label: float) -> torch.Tensor:
    loss = torch.empty((), device=pred.device, dtype=torch.float32)
    ws = _scratch(pred.device, _LOSS_WS_BYTES)
    check(load().gsb_gan_loss(ptr(pred), pred.numel(), mode, float(label), ptr(loss), None, None, ptr(ws), _LOSS_WS_BYTES,
                              _s(pred)), "gan_loss")
    return loss


def gan_loss_bwd(pred: torch.Tensor, mode: int, label: float, gscale: torch.Tensor) -> torch.Tensor:
    d = torch.empty_like(pred)
    check(load().gsb_gan_loss(ptr(pred), pred.numel(), mode, float(label), None, ptr(gscale), ptr(d), None, 0, _s(pred)),
          "gan_loss")
    return d


def l1_loss_fwd(a: torch.Tensor, b: torch.Tensor) -> torch.Tensor:
    loss = torch.empty((), device=a.device, dtype=torch.float32)
    ws = _scratch(a.device, _LOSS_WS_BYTES)
    check(load().gsb_l1_loss(ptr(a), ptr(b), a.numel(), ptr(loss), None, None, ptr(ws), _LOSS_WS_BYTES, _s(a)), "l1_loss")
    return loss


def l1_loss_bwd(a: torch.Tensor, b: torch.Tensor, gscale: torch.Tensor) -> torch.Tensor:
    d = torch.empty_like(a)
    check(load().gsb_l1_loss(ptr(a), ptr(b), a.numel(), None, ptr(gscale), ptr(d), None, 0, _s(a)), "l1_loss")
    return d


def act_bwd(out: torch.Tensor, dout: torch.Tensor, act: int) -> torch.Tensor:
    """dx = dout * act'(.) evaluated from the activation OUTPUT (contiguous fp32)."""
    dx = torch.empty_like(out)
    check(load().gsb_act_bwd(ptr(out), ptr(dout), act, ptr(dx), out.numel(), _s(out)), "act_bwd")
    return dx


# ---- optimisers -----------------------------------------------------------------------------

def rmsprop_step(p, g, square_avg, mom_buf, lr, alpha, eps, weight_decay, momentum, grad_scale=1.0) -> None:
    check(load().gsb_rmsprop_step(ptr(p), ptr(g), ptr(square_avg), ptr(mom_buf), p.numel(), lr, alpha, eps,
                                  weight_decay, momentum, grad_scale, _s(p)), "rmsprop_step")


def rmsprop_step_multi(ps, gs, sqs, bufs, lr, alpha, eps, weight_decay, momentum, grad_scale=1.0) -> None:
    if not ps:
        return
    n = _atype(C.c_int64, len(ps))(*[p.numel() for p in ps])
    check(load().gsb_rmsprop_step_multi(len(ps), _ptr_array(ps), _ptr_array(gs), _ptr_array(sqs),
                                        _ptr_array(bufs) if momentum > 0 else None, n, lr, alpha, eps, weight_decay, momentum,
                                        grad_scale, _s(ps[0])), "rmsprop_step_multi")


def adam_step_multi(ps, gs, ms, vs, lr, beta1, beta2, eps, weight_decay, step, grad_scale=1.0) -> None:
    if not ps:
        return
    n = _atype(C.c_int64, len(ps))(*[p.numel() for p in ps])
    check(load().gsb_adam_step_multi(len(ps), _ptr_array(ps), _ptr_array(gs), _ptr_array(ms), _ptr_array(vs), n, lr, beta1,
                                     beta2, eps, weight_decay, int(step), grad_scale, _s(ps[0])), "adam_step_multi")


def adam_step_multi_dev(ps, gs, ms, vs, lr, beta1, beta2, eps, weight_decay, step_dev: torch.Tensor, grad_scale=1.0) -> None:
    """Adam with the step count in device memory (int64 scalar tensor, already advanced for this step): capturable."""
    if not ps:
        return
    n = _atype(C.c_int64, len(ps))(*[p.numel() for p in ps])
    check(load().gsb_adam_step_multi_dev(len(ps), _ptr_array(ps), _ptr_array(gs), _ptr_array(ms), _ptr_array(vs), n, lr, beta1,
                                         beta2, eps, weight_decay, ptr(step_dev), grad_scale, _s(ps[0])), "adam_step_multi_dev")


def counter_add(ctr: torch.Tensor, delta: int = 1) -> None:
    """ctr (int64 scalar on the device) += delta, on the stream."""
    assert ctr.dtype == torch.int64 and ctr.is_cuda
    check(load().gsb_counter_add(ptr(ctr), int(delta), _s(ctr)), "counter_add")


def dropout_mask(shape, device, seed_k: int, calls: torch.Tensor, layer_term: int, drop_offset: int) -> torch.Tensor:
    """uint8 keep mask [n, h, w, c] with the Philox bits of seed (seed_k + calls) mod 2^62 + 1 + layer_term (calls: int64
    device scalar)."""
    mask = torch.empty(tuple(shape), device=device, dtype=torch.uint8)
    check(load().gsb_dropout_mask(ptr(mask), mask.numel(), int(seed_k) & ((1 << 64) - 1), ptr(calls), int(layer_term),
                                  int(drop_offset), _s(mask)), "dropout_mask")
    return mask


# --- device-side image chain of the end-to-end step (csrc/augment.cu) -----------------------------------
def minmax_f32(x: torch.Tensor) -> torch.Tensor:
    """Ordered keys of (min, max) over a contiguous fp32 CUDA tensor -> int32[2] on the device (opaque; feed to quantize_hist)."""
    keys = torch.empty(2, device=x.device, dtype=torch.int32)
    check(load().gsb_minmax_f32(ptr(x), x.numel(), ptr(keys), _s(x)), "minmax_f32")
    return keys


def quantize_hist(x: torch.Tensor, keys: Optional[torch.Tensor], planes: int, plane: int):
    q = torch.empty(x.shape, device=x.device, dtype=torch.uint8)
    hist = torch.empty((planes, 256), device=x.device, dtype=torch.int32)
    check(load().gsb_quantize_hist(ptr(x), ptr(keys), ptr(q), ptr(hist), planes, plane, _s(x)), "quantize_hist")
    return q, hist


def equalize_lut(q: torch.Tensor, hist: Optional[torch.Tensor], post: Optional[torch.Tensor], out: torch.Tensor,
                 planes: int, plane: int, do_equalize: bool) -> None:
    ws = None if hist is not None else torch.empty((planes, 256), device=q.device, dtype=torch.int32)
    out_f = out if out.dtype == torch.float32 else None
    out_u8 = out if out.dtype == torch.uint8 else None
    check(load().gsb_equalize_lut(ptr(q), ptr(hist), ptr(ws), ptr(post), ptr(out_f), ptr(out_u8), planes, plane,
                                  1 if do_equalize else 0, _s(q)), "equalize_lut")


def mask_affine(masks: torch.Tensor, mats: torch.Tensor, out: torch.Tensor, n: int, h: int, w: int, threshold: float) -> None:
    check(load().gsb_mask_affine(ptr(masks), ptr(mats), ptr(out), n, h, w, float(threshold), _s(masks)), "mask_affine")


# --- NVLink peer-memory all-reduce (csrc/peer.cu) ------------------------------------------------------
def peer_create():
    """-> (region pointer, IPC handle bytes) for this rank's region on the current device."""
    lib = load()
    region = C.c_void_p()
    handle = C.create_string_buffer(lib.gsb_peer_handle_bytes())
    check(lib.gsb_peer_create(C.byref(region), handle), "peer_create")
    return int(region.value), bytes(handle.raw)


def peer_open(handle: bytes) -> int:
    out = C.c_void_p()
    check(load().gsb_peer_open(C.create_string_buffer(handle, len(handle)), C.byref(out)), "peer_open")
    return int(out.value)


def peer_close(region: int) -> None:
    check(load().gsb_peer_close(C.c_void_p(region)), "peer_close")


def peer_destroy(region: int) -> None:
    check(load().gsb_peer_destroy(C.c_void_p(region)), "peer_destroy")


def peer_error(region: int) -> int:
    """0, or the sequence number of an exchange of this rank that gave up waiting for a peer (synchronises the device)."""
    seq = C.c_ulonglong(0)
    check(load().gsb_peer_error(C.c_void_p(region), C.byref(seq)), "peer_error")
    return int(seq.value)


def peer_max_elems() -> int:
    return int(load().gsb_peer_max_elems())


def peer_allreduce_f64(buf: torch.Tensor, regions, rank: int, world: int) -> None:
    """In-place sum of a contiguous fp64 device vector over the ranks whose regions are given in rank order."""
    assert buf.dtype == torch.float64 and buf.is_contiguous() and buf.is_cuda
    arr = _atype(C.c_void_p, world)(*[C.c_void_p(r) for r in regions])
    check(load().gsb_peer_allreduce_f64(ptr(buf), buf.numel(), arr, rank, world, _s(buf)), "peer_allreduce_f64")


def adam_step(p, g, exp_avg, exp_avg_sq, lr, beta1, beta2, eps, weight_decay, step, grad_scale=1.0) -> None:
    check(load().gsb_adam_step(ptr(p), ptr(g), ptr(exp_avg), ptr(exp_avg_sq), p.numel(), lr, beta1, beta2, eps,
                               weight_decay, int(step), grad_scale, _s(p)), "adam_step")
