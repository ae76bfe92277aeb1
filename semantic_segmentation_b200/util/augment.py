"""Device-side image chain of the GenSeg end-to-end step.

Mirrors running_files/train_end2end_jsrt.py:186-200 without the host round trip:

    fake_mask  = augment_masks(true_masks, draw_affine(...))          # imgaug `seq` (:100-112, :187-190) + `> 0.1` (:193)
    fake_image = netG(fake_mask)
    fake_image = synth_image_postprocess(fake_image)                  # :197-200

``synth_image_postprocess`` is bit-exact against torch / torchvision on the CPU (batch min-max normalise, uint8
quantise, per-image histogram equalisation, gamma 0.5, / 255).  The mask warp restates the imgaug chain as ONE composed
affine map per image (flip, crop-and-pad as a scale + shift, scale, translate, rotate, shear; order drawn per image)
sampled bilinearly with zero fill; imgaug itself is not available offline, so that part is pinned only by the CPU
restatement used in tests/test_augment.py ("parity unpinned" against imgaug's own resampling).
"""
from __future__ import annotations

import math
from typing import Optional

import torch

from .. import ops

_GAMMA_LUTS = {}


def gamma_lut(gamma: float, gain: float = 1.0, device=None) -> torch.Tensor:
    """256-entry uint8 LUT of torchvision.transforms.functional.adjust_gamma for uint8 images, computed with torch's
    own CPU arithmetic (x / 255 -> gain * x ** gamma -> clamp -> * 255.999 -> truncate)."""
    key = (float(gamma), float(gain), str(device))
    if key not in _GAMMA_LUTS:
        v = torch.arange(256, dtype=torch.uint8).to(torch.float32) / 255.0
        r = (gain * v ** gamma).clamp(0, 1)
        _GAMMA_LUTS[key] = r.mul(255.0 + 1.0 - 1e-3).to(torch.uint8).to(device)
    return _GAMMA_LUTS[key]


def equalize_gamma_u8(img: torch.Tensor, gamma: Optional[float] = 0.5, equalize: bool = True,
                      as_float: bool = True) -> torch.Tensor:
    """adjust_gamma(equalize(img), gamma) (/ 255.0 when ``as_float``) for a uint8 CUDA image batch [..., H, W];
    every leading index is one plane (torchvision equalises per image and channel)."""
    if img.dtype != torch.uint8 or not img.is_cuda:
        raise RuntimeError("equalize_gamma_u8: expected a uint8 CUDA tensor")
    q = img.contiguous()
    plane = q.shape[-1] * q.shape[-2]
    planes = q.numel() // plane
    post = gamma_lut(gamma, device=q.device) if gamma is not None else None
    out = torch.empty(q.shape, device=q.device, dtype=torch.float32 if as_float else torch.uint8)
    ops.equalize_lut(q, None, post, out, planes, plane, equalize)
    return out


def synth_image_postprocess(fake_image: torch.Tensor, gamma: float = 0.5) -> torch.Tensor:
    """train_end2end_jsrt.py:197-200: ((x - min) / (max - min)) over the whole batch -> uint8 -> equalize -> gamma -> / 255."""
    if not fake_image.is_cuda:
        raise RuntimeError("synth_image_postprocess: expected a CUDA tensor (no CPU fallback)")
    x = fake_image.detach().float().contiguous()
    plane = x.shape[-1] * x.shape[-2]
    planes = x.numel() // plane
    keys = ops.minmax_f32(x)
    q, hist = ops.quantize_hist(x, keys, planes, plane)
    out = torch.empty_like(x)
    ops.equalize_lut(q, hist, gamma_lut(gamma, device=x.device), out, planes, plane, True)
    return out


# --- mask augmentation -------------------------------------------------------------------------------
def _mat(a, b, c, d, e, f):
    return [[a, b, c], [d, e, f], [0.0, 0.0, 1.0]]


def _mul(p, q):
    return [[sum(p[i][k] * q[k][j] for k in range(3)) for j in range(3)] for i in range(3)]


def _about_centre(m, h, w):
    cx, cy = (w - 1) / 2.0, (h - 1) / 2.0
    return _mul(_mat(1, 0, cx, 0, 1, cy), _mul(m, _mat(1, 0, -cx, 0, 1, -cy)))


def draw_affine(n: int, h: int, w: int, generator: Optional[torch.Generator] = None) -> torch.Tensor:
    """Per-image OUTPUT->INPUT pixel maps [n, 6] of the reference's augmenter (train_end2end_jsrt.py:100-112):
    Fliplr(0.5), CropAndPad(percent=(0, 0.1)) (per-side pad fraction, image resized back), Affine scale (0.8, 1.2) per axis,
    translate_percent (-0.1, 0.1) per axis, rotate (-15, 15) deg, shear (-8, 8) deg, applied in a random order."""
    g = generator
    u = torch.rand(n, 12, generator=g).tolist()
    perms = [torch.randperm(6, generator=g).tolist() for _ in range(n)]
    out = []
    for i in range(n):
        r = u[i]
        steps = []
        # forward (input -> output) maps
        steps.append(_mat(-1, 0, w - 1, 0, 1, 0) if r[0] < 0.5 else _mat(1, 0, 0, 0, 1, 0))                 # Fliplr
        top, right, bottom, left = (0.1 * r[1] * h, 0.1 * r[2] * w, 0.1 * r[3] * h, 0.1 * r[4] * w)          # CropAndPad
        sx, sy = w / (w + left + right), h / (h + top + bottom)
        steps.append(_mat(sx, 0, left * sx, 0, sy, top * sy))
        steps.append(_about_centre(_mat(0.8 + 0.4 * r[5], 0, 0, 0, 0.8 + 0.4 * r[6], 0), h, w))             # scale
        steps.append(_mat(1, 0, (-0.1 + 0.2 * r[7]) * w, 0, 1, (-0.1 + 0.2 * r[8]) * h))                   # translate
        a = math.radians(-15.0 + 30.0 * r[9])
        steps.append(_about_centre(_mat(math.cos(a), -math.sin(a), 0, math.sin(a), math.cos(a), 0), h, w))  # rotate
        sh = math.radians(-8.0 + 16.0 * r[10])
        steps.append(_about_centre(_mat(1, -math.tan(sh), 0, 0, 1, 0), h, w))                                # shear
        fwd = _mat(1, 0, 0, 0, 1, 0)
        for k in perms[i]:
            fwd = _mul(steps[k], fwd)
        inv = torch.linalg.inv(torch.tensor(fwd, dtype=torch.float64))
        out.append(inv[:2].reshape(6))
    return torch.stack(out).to(torch.float32)


def augment_masks(masks: torch.Tensor, mats: torch.Tensor, threshold: float = 0.1) -> torch.Tensor:
    """where(warp(mask) > threshold, 1, 0) for float masks [N, 1, H, W] or [N, H, W] on the GPU (:187-195)."""
    if not masks.is_cuda:
        raise RuntimeError("augment_masks: expected a CUDA tensor (no CPU fallback)")
    m = masks.detach().float().contiguous()
    h, w = m.shape[-2], m.shape[-1]
    n = m.numel() // (h * w)
    if mats.shape != (n, 6):
        raise RuntimeError("augment_masks: mats must be [%d, 6]" % n)
    out = torch.empty_like(m)
    ops.mask_affine(m, mats.to(device=m.device, dtype=torch.float32).contiguous(), out, n, h, w, threshold)
    return out
