"""TEST INFRASTRUCTURE ONLY (imported by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg).

CPU restatement of the image chain of the GenSeg end-to-end step, running_files/train_end2end_jsrt.py:186-200:

  synth_image_postprocess : lines :197-200 verbatim on the CPU -- torch min/max, mul/add_/clamp_/to(uint8), then
                            torchvision.transforms.functional.equalize and adjust_gamma(.., 0.5), / 255.0.
                            torchvision (pinned 0.14.1 in env.sh:3; 0.26.0 here) is the third-party owner of the
                            equalise / gamma arithmetic; `equalize_plane` below also restates its published algorithm
                            (_functional_tensor._scale_channel) in numpy so the fixture does not depend on its version.
  warp_masks              : the imgaug chain (:100-112, :187-190) restated as one inverse-mapped bilinear warp per image
                            with zero fill, then `> 0.1` (:193-195).  imgaug is absent offline: PARITY UNPINNED against
                            imgaug's own resampling; pinned only against this restatement.
"""
from __future__ import annotations

import numpy as np
import torch


def equalize_plane(q: np.ndarray) -> np.ndarray:
    """torchvision _scale_channel on one uint8 plane."""
    hist = np.bincount(q.reshape(-1), minlength=256).astype(np.int64)
    nz = hist[hist != 0]
    step = int(nz[:-1].sum()) // 255
    if step == 0:
        return q.copy()
    lut = (np.cumsum(hist) + step // 2) // step
    lut = np.clip(np.concatenate([[0], lut[:-1]]), 0, 255).astype(np.uint8)
    return lut[q]


def gamma_u8(q: np.ndarray, gamma: float = 0.5) -> np.ndarray:
    """torchvision adjust_gamma on uint8: x/255 -> x**gamma -> clamp -> * 255.999 -> truncate (fp32)."""
    r = (torch.from_numpy(q).to(torch.float32) / 255.0) ** gamma
    return r.clamp(0, 1).mul(255.0 + 1.0 - 1e-3).to(torch.uint8).numpy()


def synth_image_postprocess(fake_image: torch.Tensor, gamma: float = 0.5, use_torchvision: bool = True) -> torch.Tensor:
    x = fake_image.detach().float().cpu()
    x = (x - x.min()) / (x.max() - x.min())
    q = x.mul(255).add_(0.5).clamp_(0, 255).to(torch.uint8)
    if use_torchvision:
        import torchvision.transforms.functional as TF
        return TF.adjust_gamma(TF.equalize(q), gamma) / 255.0
    planes = q.reshape(-1, q.shape[-2], q.shape[-1]).numpy()
    out = np.stack([gamma_u8(equalize_plane(p), gamma) for p in planes]).reshape(q.shape)
    return torch.from_numpy(out) / 255.0


def warp_masks(masks: torch.Tensor, mats: torch.Tensor, threshold: float = 0.1) -> torch.Tensor:
    """fp32 inverse-mapped bilinear warp, zero outside, same operation order as csrc/augment.cu (no FMA on the CPU)."""
    m = masks.detach().float().cpu().numpy()
    shape = m.shape
    h, w = shape[-2], shape[-1]
    m = m.reshape(-1, h, w)
    mats = mats.detach().float().cpu().numpy().astype(np.float32)
    xo, yo = np.meshgrid(np.arange(w, dtype=np.float32), np.arange(h, dtype=np.float32))
    out = np.empty_like(m)
    f = np.float32
    for n in range(m.shape[0]):
        a = mats[n]
        xs = (a[0] * xo + a[1] * yo) + a[2]
        ys = (a[3] * xo + a[4] * yo) + a[5]
        xf, yf = np.floor(xs), np.floor(ys)
        ax, ay = xs - xf, ys - yf
        x0, y0 = xf.astype(np.int64), yf.astype(np.int64)
        pad = np.zeros((h + 2, w + 2), dtype=np.float32)
        pad[1:-1, 1:-1] = m[n]

        def at(yy, xx):
            ok = (yy >= 0) & (yy < h) & (xx >= 0) & (xx < w)
            return np.where(ok, pad[np.clip(yy, -1, h) + 1, np.clip(xx, -1, w) + 1], f(0))

        top = at(y0, x0) * (f(1) - ax) + at(y0, x0 + 1) * ax
        bot = at(y0 + 1, x0) * (f(1) - ax) + at(y0 + 1, x0 + 1) * ax
        v = top * (f(1) - ay) + bot * ay
        out[n] = v if threshold < 0 else (v > f(threshold)).astype(np.float32)
    return torch.from_numpy(out.reshape(shape))
