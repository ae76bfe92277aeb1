"""ctypes binding of libgenseg_b200.so (include/genseg_b200.h).

The product path has NO fallback: if the shared library is missing or a call fails, a
RuntimeError is raised.  PyTorch is used for device memory, streams and autograd plumbing only.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Optional

import torch

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libgenseg_b200.so")

GSB_F32, GSB_BF16 = 0, 1
ACT_NONE, ACT_RELU, ACT_LRELU, ACT_TANH = 0, 1, 2, 3
MODE_CONV, MODE_CONVT = 0, 1


class GsbView(C.Structure):
    _fields_ = [("ptr", C.c_void_p), ("dtype", C.c_int32), ("n", C.c_int32), ("h", C.c_int32), ("w", C.c_int32),
                ("c", C.c_int32), ("sn", C.c_int64), ("sh", C.c_int64), ("sw", C.c_int64), ("sc", C.c_int64)]


class GsbConvGeom(C.Structure):
    _fields_ = [("mode", C.c_int32), ("kh", C.c_int32), ("kw", C.c_int32), ("stride", C.c_int32), ("pad", C.c_int32)]


_VP = C.POINTER(GsbView)
_GP = C.POINTER(GsbConvGeom)
_p, _i, _i64, _f, _d, _u64, _sz = C.c_void_p, C.c_int, C.c_int64, C.c_float, C.c_double, C.c_uint64, C.c_size_t

# name -> (restype, argtypes); every symbol declared in include/genseg_b200.h
SIGNATURES = {
    "gsb_version": (_i, []),
    "gsb_last_error": (C.c_char_p, []),
    "gsb_set_sm_reserve": (_i, [_i]),
    "gsb_tc_available": (_i, []),
    "gsb_launch_count": (C.c_ulonglong, []),
    "gsb_copy_view": (_i, [_VP, _VP, _p]),
    "gsb_cast_flat": (_i, [_p, _i, _p, _i, _i64, _p]),
    "gsb_pack_weight": (_i, [_p, _i, _i, _i, _i, _i, _p, _i, _p]),
    "gsb_unpack_wgrad": (_i, [_p, _i, _i, _i, _i, _i, _i, _p, _i, _p]),
    "gsb_sum_slices": (_i, [_p, _i, _i64, _p, _i, _p]),
    "gsb_pack_weight_multi": (_i, [_i, _p, _p, _p, _p, _p, _p, _i, _p]),
    "gsb_unpack_wgrad_multi": (_i, [_i, _p, _p, _p, _p, _p, _p, _p, _i, _p]),
    "gsb_merge_upconv": (_i, [_p, _p, _p, _p, _p, _p, _p, _i, _i, _p, _p, _p]),
    "gsb_unmerge_upconv": (_i, [_p] * 9 + [_i, _i] + [_p] * 7 + [_p, _sz, _p]),
    "gsb_conv_gather": (_i, [_VP, _p, _i, _p, _VP, _GP, _i, _i, _i, _p, _p, _p, _sz, _p]),
    "gsb_conv_gather_affine": (_i, [_VP, _p, _i, _p, _p, _VP, _GP, _i, _p]),
    "gsb_conv_stats_rows": (_sz, [_VP, _GP]),
    "gsb_conv_wgrad_plan": (_i, [_VP, _VP, _GP, _i]),
    "gsb_conv_wgrad": (_i, [_VP, _VP, _p, _i, _p, _GP, _i, _p]),
    "gsb_edge_rows": (_i, []),
    "gsb_stem_stats_rows": (_i, []),
    "gsb_stem_conv_fwd": (_i, [_p, _i, _i, _i, _i, _p, _p, _VP, _i, _i, _i, _i, _p, _p, _p]),
    "gsb_stem_conv_fwd_affine": (_i, [_p, _i, _i, _i, _i, _p, _p, _p, _VP, _i, _i, _i, _i, _p]),
    "gsb_stem_conv_wgrad": (_i, [_p, _i, _i, _i, _i, _VP, _i, _i, _i, _p, _p, _sz, _p]),
    "gsb_head_conv_fwd": (_i, [_VP, _p, _p, _i, _p, _p]),
    "gsb_head_conv_bwd": (_i, [_VP, _p, _i, _p, _VP, _p, _p, _p, _sz, _p]),
    "gsb_bias_grad_workspace_bytes": (_sz, [_VP]),
    "gsb_bias_grad": (_i, [_VP, _p, _p, _sz, _p]),
    "gsb_bn_workspace_bytes": (_sz, [_VP]),
    "gsb_bn_reduce": (_i, [_VP, _i, _p, _p, _sz, _p]),
    "gsb_bn_reduce_partials": (_i, [_p, _i, _i, _p, _p]),
    "gsb_bn_finalize": (_i, [_p, _d, _p, _p, _p, _p, _p, _f, _f, _i, _p, _p, _p, _p, _i, _p]),
    "gsb_bn_act": (_i, [_VP, _p, _p, _p, _i, _VP, _i, _VP, _u64, _i64, _p, _i, _p]),
    "gsb_bn_act_pool": (_i, [_VP, _p, _p, _VP, _VP, _p]),
    "gsb_bn_bwd_reduce": (_i, [_VP, _VP, _VP, _p, _p, _p, _p, _p, _i, _i, _u64, _i64, _p, _p, _p, _sz, _p]),
    "gsb_bn_bwd_apply": (_i, [_VP, _VP, _VP, _p, _p, _p, _p, _p, _i, _i, _u64, _i64, _p, _p, _d, _VP, _p, _p, _i, _p]),
    "gsb_bn_param_grad": (_i, [_p, _i, _p, _p, _i, _p]),
    "gsb_bn_stats_finalize": (_i, [_p, _i, _i, _d, _p, _i, _i, _p, _p, _p, _p, _p, _f, _f, _p, _p, _p, _p, _i, _p]),
    "gsb_bn_bwd_partial": (_i, [_VP, _VP, _VP, _p, _p, _p, _p, _p, _i, _i, _u64, _i64, _p, _p, _sz, _p, _i, _p]),
    "gsb_bn_bwd_partial_pool": (_i, [_VP, _VP, _VP, _p, _p, _p, _p, _p, _sz, _p, _p]),
    "gsb_bn_bwd_stats": (_i, [_p, _i, _i, _p, _i, _i, _p, _p, _i, _p, _p]),
    "gsb_maxpool2_fwd": (_i, [_VP, _VP, _p]),
    "gsb_maxpool2_bwd": (_i, [_VP, _VP, _VP, _i, _p]),
    "gsb_upsample2_fwd": (_i, [_VP, _VP, _i, _i, _p]),
    "gsb_upsample2_bwd": (_i, [_VP, _VP, _i, _i, _p]),
    "gsb_seg_loss_reduce": (_i, [_p, _p, _i, _i, _i, _i, _p, _p, _sz, _p]),
    "gsb_seg_loss_workspace_bytes": (_sz, [_i, _i, _i, _i]),
    "gsb_seg_loss_finalize": (_i, [_p, _d, _p, _p]),
    "gsb_seg_loss_bwd": (_i, [_p, _p, _i, _i, _i, _i, _p, _d, _p, _p, _p]),
    "gsb_eval_counts": (_i, [_p, _p, _i, _i, _i, _i, _p, _p, _p]),
    "gsb_dice_sums": (_i, [_p, _p, _i64, _i64, _p, _p, _sz, _p]),
    "gsb_dice_sums_noalloc": (_i, [_p, _p, _i64, _i64, _p, _p]),
    "gsb_dice_sums_noalloc_max_slices": (_sz, []),
    "gsb_gan_loss": (_i, [_p, _i64, _i, _f, _p, _p, _p, _p, _sz, _p]),
    "gsb_l1_loss": (_i, [_p, _p, _i64, _p, _p, _p, _p, _sz, _p]),
    "gsb_act_bwd": (_i, [_p, _p, _i, _p, _i64, _p]),
    "gsb_rmsprop_step": (_i, [_p, _p, _p, _p, _i64, _f, _f, _f, _f, _f, _f, _p]),
    "gsb_adam_step": (_i, [_p, _p, _p, _p, _i64, _f, _f, _f, _f, _f, _i64, _f, _p]),
    "gsb_rmsprop_step_multi": (_i, [_i, _p, _p, _p, _p, _p, _f, _f, _f, _f, _f, _f, _p]),
    "gsb_adam_step_multi": (_i, [_i, _p, _p, _p, _p, _p, _f, _f, _f, _f, _f, _i64, _f, _p]),
    "gsb_adam_step_multi_dev": (_i, [_i, _p, _p, _p, _p, _p, _f, _f, _f, _f, _f, _p, _f, _p]),
    "gsb_counter_add": (_i, [_p, _i64, _p]),
    "gsb_dropout_mask": (_i, [_p, _i64, _u64, _p, _u64, _i64, _p]),
    "gsb_minmax_f32": (_i, [_p, _i64, _p, _p]),
    "gsb_quantize_hist": (_i, [_p, _p, _p, _p, _i, _i, _p]),
    "gsb_equalize_lut": (_i, [_p, _p, _p, _p, _p, _p, _i, _i, _i, _p]),
    "gsb_mask_affine": (_i, [_p, _p, _p, _i, _i, _i, _f, _p]),
    "gsb_peer_region_bytes": (_i64, []),
    "gsb_peer_max_elems": (_i, []),
    "gsb_peer_handle_bytes": (_i, []),
    "gsb_peer_create": (_i, [_p, _p]),
    "gsb_peer_open": (_i, [_p, _p]),
    "gsb_peer_close": (_i, [_p]),
    "gsb_peer_destroy": (_i, [_p]),
    "gsb_peer_error": (_i, [_p, _p]),
    "gsb_peer_allreduce_f64": (_i, [_p, _i, _p, _i, _i, _p]),
}

_lib: Optional[C.CDLL] = None


def load() -> C.CDLL:
    """Load (once) the in-tree shared library; raise loudly when it is missing."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            "libgenseg_b200.so is not built (%s).  Run `python -m semantic_segmentation_b200.build` "
            "(or __graft_entry__.build()).  There is no CPU / PyTorch fallback." % LIB_PATH)
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def check(rc: int, what: str = "") -> None:
    if rc != 0:
        msg = load().gsb_last_error().decode("utf-8", "replace")
        raise RuntimeError("genseg_b200 %s failed (%d): %s" % (what, rc, msg))


_RAW_STREAM = getattr(torch._C, "_cuda_getCurrentRawStream", None)


def stream_ptr(device=None) -> int:
    """cudaStream_t of torch's current stream on ``device`` (honours ``torch.cuda.stream(...)`` and graph capture).
    ``torch.cuda.current_stream()`` builds a Stream object per call (~7 us, 16 % of the host time of an eager Pix2Pix step
    with its 510 launches); the raw accessor returns the handle directly."""
    if _RAW_STREAM is not None:
        idx = device if isinstance(device, int) else getattr(device, "index", None)
        if idx is None:
            idx = torch.cuda.current_device()
        return _RAW_STREAM(idx)
    return torch.cuda.current_stream(device).cuda_stream


def dtype_code(t: torch.dtype) -> int:
    if t == torch.float32:
        return GSB_F32
    if t == torch.bfloat16:
        return GSB_BF16
    raise RuntimeError("genseg_b200: unsupported dtype %s" % t)


def require_cuda(t: torch.Tensor, name: str = "tensor") -> None:
    if not t.is_cuda:
        raise RuntimeError("genseg_b200: %s must be a CUDA tensor (the hot path has no CPU fallback)" % name)


def _canon(n, h, w, c, sn, sh, sw, sc):
    """Strides of size-1 dimensions are arbitrary in torch; replace them by the values that keep the view
    pixel-linear (sh = w*sw, sn = h*sh) so the kernels' layout checks see a canonical description."""
    if c == 1:
        sc = 1
    if w == 1:
        sw = sh if h > 1 else (sn if n > 1 else c * sc)
    if h == 1:
        sh = w * sw
    if n == 1:
        sn = h * sh
    return sn, sh, sw, sc


def view_nhwc(t: torch.Tensor) -> GsbView:
    """View of a [N,H,W,C]-shaped tensor (any strides: NHWC buffers and their channel slices)."""
    require_cuda(t)
    n, h, w, c = t.shape
    sn, sh, sw, sc = _canon(n, h, w, c, *t.stride())
    return GsbView(t.data_ptr(), dtype_code(t.dtype), n, h, w, c, sn, sh, sw, sc)


def view_nchw(t: torch.Tensor) -> GsbView:
    """View of a [N,C,H,W]-shaped tensor as (n,h,w,c) with its own strides."""
    require_cuda(t)
    n, c, h, w = t.shape
    sn, sc, sh, sw = t.stride()
    sn, sh, sw, sc = _canon(n, h, w, c, sn, sh, sw, sc)
    return GsbView(t.data_ptr(), dtype_code(t.dtype), n, h, w, c, sn, sh, sw, sc)


def ptr(t: Optional[torch.Tensor]) -> Optional[int]:
    if t is None:
        return None
    require_cuda(t)
    return t.data_ptr()
