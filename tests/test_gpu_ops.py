"""GPU parity tests of the individual C-ABI ops against the same torch CPU functionals the oracle uses."""
import copy
import math

import numpy as np
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu

from semantic_segmentation_b200 import ops  # noqa: E402
from semantic_segmentation_b200._lib import (ACT_LRELU, ACT_NONE, ACT_RELU, ACT_TANH, MODE_CONV, MODE_CONVT,  # noqa: E402
                                              view_nchw, view_nhwc)

DEV = "cuda"


def rel(a, b):
    a = a.detach().double().cpu().flatten()
    b = b.detach().double().cpu().flatten()
    return float((a - b).norm() / b.norm().clamp_min(1e-30))


def nhwc(x, dtype):
    return x.permute(0, 2, 3, 1).contiguous().to(DEV, dtype)


def from_nhwc(x):
    return x.float().cpu().permute(0, 3, 1, 2).contiguous()


GEOMS = [  # kind, k, stride, pad, cin, cout, n, h, w
    ("conv", 3, 1, 1, 64, 64, 2, 16, 16),
    ("conv", 3, 1, 1, 128, 64, 1, 24, 40),
    ("conv", 3, 1, 1, 64, 128, 3, 9, 11),
    ("conv", 4, 2, 1, 64, 128, 2, 32, 32),
    ("conv", 4, 1, 1, 128, 64, 2, 31, 31),
    ("conv", 1, 1, 0, 64, 64, 2, 8, 8),
    ("convT", 2, 2, 0, 128, 64, 2, 8, 8),
    ("convT", 4, 2, 1, 64, 64, 2, 8, 8),
    ("convT", 8, 2, 3, 128, 64, 2, 4, 4),
    ("convT", 8, 2, 3, 64, 64, 3, 1, 1),
    ("conv", 4, 2, 1, 64, 64, 5, 2, 2),
    # halo-stationary tcgen05 kernel (tiled space >= 16x16): full / partial tiles, 2 N-tiles, 4x4 windows, parity phases
    ("conv", 3, 1, 1, 128, 128, 2, 32, 48),
    ("conv", 3, 1, 1, 64, 256, 1, 16, 16),
    ("conv", 3, 1, 1, 192, 64, 1, 20, 17),
    ("convT", 8, 2, 3, 64, 64, 1, 16, 16),
    ("convT", 2, 2, 0, 128, 64, 2, 16, 16),
    ("conv", 4, 2, 1, 128, 128, 1, 64, 64),
]


def conv_case(kind, k, stride, pad, cin, cout, n, h, w, dtype, use_tc, seed=0):
    g = torch.Generator().manual_seed(seed)
    x = torch.randn(n, cin, h, w, generator=g)
    wshape = (cout, cin, k, k) if kind == "conv" else (cin, cout, k, k)
    wt = torch.randn(wshape, generator=g) / math.sqrt(cin * k * k)
    if dtype == torch.bfloat16:
        x = x.bfloat16().float()
        wt = wt.bfloat16().float()
    x.requires_grad_(True)
    wt.requires_grad_(True)
    y = F.conv2d(x, wt, None, stride=stride, padding=pad) if kind == "conv" else \
        F.conv_transpose2d(x, wt, None, stride=stride, padding=pad)
    dy = torch.randn(y.shape, generator=g)
    if dtype == torch.bfloat16:
        dy = dy.bfloat16().float()
    y.backward(dy)
    fwd_mode = MODE_CONV if kind == "conv" else MODE_CONVT
    bwd_mode = MODE_CONVT if kind == "conv" else MODE_CONV
    xd = nhwc(x.detach(), dtype)
    wd = wt.detach().to(DEV)
    wp = ops.pack_weight(wd, kind == "conv", dtype)
    wpb = ops.pack_weight(wd, kind != "conv", dtype)
    oh, ow = y.shape[2], y.shape[3]
    yd = torch.empty((n, oh, ow, cout), device=DEV, dtype=dtype)
    stats, rows = ops.conv_gather(view_nhwc(xd), wp, None, view_nhwc(yd), fwd_mode, k, stride, pad, use_tc=use_tc,
                                  want_stats=(use_tc == 1 and not (kind == "convT" and stride > 1)))
    dyd = nhwc(dy, dtype)
    dxd = torch.empty_like(xd)
    ops.conv_gather(view_nhwc(dyd), wpb, None, view_nhwc(dxd), bwd_mode, k, stride, pad, use_tc=use_tc)
    dwp = torch.zeros((k * k, cout, cin), device=DEV, dtype=torch.float32)
    ops.conv_wgrad_into(view_nhwc(xd), view_nhwc(dyd), dwp, fwd_mode, k, stride, pad, use_tc=use_tc)
    dw = ops.unpack_wgrad(dwp, wshape, kind == "conv")
    torch.cuda.synchronize()
    res = dict(fwd=rel(from_nhwc(yd), y), dgrad=rel(from_nhwc(dxd), x.grad), wgrad=rel(dw, wt.grad))
    if stats is not None:
        sums = ops.bn_reduce_partials(stats, rows, cout).cpu()
        yq = from_nhwc(yd).double()
        ref = torch.cat([yq.sum((0, 2, 3)), (yq * yq).sum((0, 2, 3))])
        res["stats"] = rel(sums, ref)
    return res


@pytest.mark.parametrize("geom", GEOMS, ids=[f"{g[0]}k{g[1]}s{g[2]}_{g[4]}to{g[5]}_{g[7]}x{g[8]}" for g in GEOMS])
def test_conv_direct_fp32(geom):
    res = conv_case(*geom, dtype=torch.float32, use_tc=0)
    assert res["fwd"] < 1e-5 and res["dgrad"] < 1e-5 and res["wgrad"] < 2e-5, res


@pytest.mark.parametrize("geom", GEOMS, ids=[f"{g[0]}k{g[1]}s{g[2]}_{g[4]}to{g[5]}_{g[7]}x{g[8]}" for g in GEOMS])
def test_conv_direct_bf16(geom):
    res = conv_case(*geom, dtype=torch.bfloat16, use_tc=0)
    assert res["fwd"] < 6e-3 and res["dgrad"] < 6e-3 and res["wgrad"] < 1e-4, res


@pytest.mark.parametrize("geom", GEOMS, ids=[f"{g[0]}k{g[1]}s{g[2]}_{g[4]}to{g[5]}_{g[7]}x{g[8]}" for g in GEOMS])
def test_conv_tcgen05_bf16(geom):
    """tcgen05/TMEM/TMA path: operands are exact bf16, accumulation fp32 -> only the bf16 rounding of the
    stored output separates it from the fp32 reference (rel-L2 <= 1e-2 per BASELINE north_star)."""
    res = conv_case(*geom, dtype=torch.bfloat16, use_tc=1)
    assert res["fwd"] < 6e-3 and res["dgrad"] < 6e-3 and res["wgrad"] < 1e-4, res
    if "stats" in res:
        assert res["stats"] < 1e-5, res


TILE_GEOMS = [("conv", 3, 1, 1, 128, 256, 3, 40, 24), ("conv", 4, 2, 1, 64, 128, 2, 32, 64), ("convT", 4, 2, 1, 128, 256, 1, 16, 24)]


@pytest.mark.parametrize("pair", ["0", "2"])
@pytest.mark.parametrize("tile", ["256,2", "256,1", "128,2", "128,1", "64,2", "64,1"])
@pytest.mark.parametrize("geom", TILE_GEOMS, ids=[f"{g[0]}k{g[1]}s{g[2]}_{g[4]}to{g[5]}_{g[7]}x{g[8]}" for g in TILE_GEOMS])
def test_conv_tcgen05_tile_shapes(geom, tile, pair):
    """Every (BN output channels, mt M-tiles) shape of the halo kernel on the same problem (the cost model would
    pick only one of them): 16- and 8-pixel-wide tiles, partial tiles in x and y, odd tile counts (ghost cluster tile);
    each with one-CTA MMAs (GSB_TC_PAIR=0) and with the CTA-pair form (tcgen05 cta_group::2, M = 256) for every width."""
    import os
    os.environ["GSB_TC_TILE"] = tile
    os.environ["GSB_TC_PAIR"] = pair
    try:
        res = conv_case(*geom, dtype=torch.bfloat16, use_tc=1, seed=3)
    finally:
        del os.environ["GSB_TC_TILE"]
        del os.environ["GSB_TC_PAIR"]
    assert res["fwd"] < 6e-3 and res["dgrad"] < 6e-3 and res["wgrad"] < 1e-4, res
    if "stats" in res:
        assert res["stats"] < 1e-5, res


# Every tensor-core convolution of the bench step (BASELINE config 2: UNet(1,2,bilinear) at 256x256), at the per-GPU batch
# of 1 rank (64 images) and of 8 ranks (8 images): the shapes whose kernel variants (BN=256 cluster path, 64-column tiles,
# one-tile-per-CTA launches, split-K weight gradients) the bench actually times.
PROD_LAYERS = [  # cin, cout, hw
    (64, 64, 256), (128, 64, 256), (64, 128, 128), (128, 128, 128), (256, 128, 128), (128, 64, 128),
    (128, 256, 64), (256, 256, 64), (512, 256, 64), (256, 128, 64), (256, 512, 32), (512, 512, 32), (1024, 512, 32),
    (512, 256, 32), (512, 512, 16),
]


def _rel_dev(a: torch.Tensor, b_cpu: torch.Tensor) -> float:
    b = b_cpu.to(DEV)
    num = (a.float() - b).double().square().sum()
    return float((num / b.double().square().sum().clamp_min(1e-300)).sqrt())


@pytest.mark.parametrize("batch", [64, 8])
@pytest.mark.parametrize("layer", PROD_LAYERS, ids=["%dto%d_%d" % l for l in PROD_LAYERS])
def test_conv_tcgen05_production_shapes(layer, batch):
    """fprop / dgrad / wgrad of one 3x3 layer at its production shape against torch's CPU fp32 convolution on the same
    bf16-representable operands: <= 6e-3 (bf16 rounding of the stored output; north_star per-layer bound 1e-2) for the
    activations and input gradients, <= 2e-4 for the fp32 weight gradient (fp32 accumulation in TMEM over the up to 113 K
    pixels of one split, against a CPU reference that also accumulates in fp32: measured 0.5e-4 .. 1.03e-4, growing with
    the pixels per split); the fused BatchNorm statistics of the epilogue <= 1e-5."""
    cin, cout, hw = layer
    g = torch.Generator().manual_seed(cin * 7 + cout * 3 + hw + batch)
    x = torch.randn(batch, cin, hw, hw, generator=g).bfloat16().float()
    wt = (torch.randn(cout, cin, 3, 3, generator=g) / math.sqrt(cin * 9)).bfloat16().float()
    dy = torch.randn(batch, cout, hw, hw, generator=g).bfloat16().float()
    x.requires_grad_(True)
    wt.requires_grad_(True)
    y = F.conv2d(x, wt, None, padding=1)
    y.backward(dy)
    xd = nhwc(x.detach(), torch.bfloat16)
    dyd = nhwc(dy, torch.bfloat16)
    wd = wt.detach().to(DEV)
    wp, wpb = ops.pack_weight(wd, True, torch.bfloat16), ops.pack_weight(wd, False, torch.bfloat16)
    yd = torch.empty((batch, hw, hw, cout), device=DEV, dtype=torch.bfloat16)
    stats, rows = ops.conv_gather(view_nhwc(xd), wp, None, view_nhwc(yd), MODE_CONV, 3, 1, 1, use_tc=1, want_stats=True)
    dxd = torch.empty_like(xd)
    ops.conv_gather(view_nhwc(dyd), wpb, None, view_nhwc(dxd), MODE_CONVT, 3, 1, 1, use_tc=1)
    dwp = torch.zeros((9, cout, cin), device=DEV, dtype=torch.float32)
    ops.conv_wgrad_into(view_nhwc(xd), view_nhwc(dyd), dwp, MODE_CONV, 3, 1, 1, use_tc=1)
    dw = ops.unpack_wgrad(dwp, (cout, cin, 3, 3), True)
    torch.cuda.synchronize()
    res = dict(fwd=_rel_dev(yd.permute(0, 3, 1, 2), y.detach()), dgrad=_rel_dev(dxd.permute(0, 3, 1, 2), x.grad),
               wgrad=_rel_dev(dw, wt.grad))
    sums = ops.bn_reduce_partials(stats, rows, cout)
    yq = yd.double()
    res["stats"] = rel(sums, torch.cat([yq.sum((0, 1, 2)), (yq * yq).sum((0, 1, 2))]))
    print(layer, batch, res)
    assert res["fwd"] < 6e-3 and res["dgrad"] < 6e-3 and res["wgrad"] < 2e-4 and res["stats"] < 1e-5, res


def test_edge_strided_views_direct():
    """NCHW fp32 network edges (Cin=1 input, Cout=2 output with bias+tanh) through the SIMT kernel."""
    g = torch.Generator().manual_seed(3)
    x = torch.rand(2, 1, 20, 28, generator=g)
    w = torch.randn(64, 1, 3, 3, generator=g)
    y = F.conv2d(x, w, None, padding=1)
    xd, wd = x.to(DEV), w.to(DEV)
    yd = torch.empty((2, 20, 28, 64), device=DEV, dtype=torch.float32)
    ops.conv_gather(view_nchw(xd), ops.pack_weight(wd, True, torch.float32), None, view_nhwc(yd), MODE_CONV, 3, 1, 1,
                    use_tc=0)
    assert rel(from_nhwc(yd), y) < 1e-5
    w2 = torch.randn(2, 64, 1, 1, generator=g) * 0.1
    b2 = torch.randn(2, generator=g)
    z = torch.tanh(F.conv2d(y, w2, b2))
    zd = torch.empty((2, 2, 20, 28), device=DEV, dtype=torch.float32)
    ops.conv_gather(view_nhwc(yd), ops.pack_weight(w2.to(DEV), True, torch.float32), b2.to(DEV), view_nchw(zd), MODE_CONV,
                    1, 1, 0, act=ACT_TANH, use_tc=0)
    assert rel(zd, z) < 1e-5


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("c", [64, 128, 1024])
def test_bn_act_fwd_bwd(dtype, c):
    g = torch.Generator().manual_seed(c)
    n, h, w = 3, 10, 6
    x = torch.randn(n, c, h, w, generator=g) * 1.5 + 0.3
    if dtype == torch.bfloat16:
        x = x.bfloat16().float()
    gamma = 1 + 0.1 * torch.randn(c, generator=g)
    beta = 0.1 * torch.randn(c, generator=g)
    rm, rv = torch.zeros(c), torch.ones(c)
    xr = x.clone().requires_grad_(True)
    gr, br = gamma.clone().requires_grad_(True), beta.clone().requires_grad_(True)
    rm_ref, rv_ref = rm.clone(), rv.clone()
    z = F.batch_norm(xr, rm_ref, rv_ref, gr, br, training=True, momentum=0.1, eps=1e-5)
    a1, a2 = F.leaky_relu(z, 0.2), F.relu(z)
    d1, d2 = torch.randn(a1.shape, generator=g), torch.randn(a1.shape, generator=g)
    if dtype == torch.bfloat16:
        d1, d2 = d1.bfloat16().float(), d2.bfloat16().float()
    (a1 * d1 + a2 * d2).sum().backward()
    xd = nhwc(x, dtype)
    sums = ops.bn_reduce(xd)
    rmd, rvd = rm.to(DEV), rv.to(DEV)
    nbt = torch.zeros((), dtype=torch.int64, device=DEV)
    scale, shift, mean, invstd = ops.bn_finalize(sums, n * h * w, gamma.to(DEV), beta.to(DEV), rmd, rvd, nbt, 0.1, 1e-5,
                                                 c, torch.device(DEV))
    o1, o2 = torch.empty_like(xd), torch.empty_like(xd)
    ops.bn_act(xd, scale, shift, ACT_LRELU, o1, ACT_RELU, o2)
    tol = 1e-5 if dtype == torch.float32 else 6e-3
    assert rel(from_nhwc(o1), a1) < tol and rel(from_nhwc(o2), a2) < tol
    assert rel(rmd, rm_ref) < 1e-5 and rel(rvd, rv_ref) < 1e-5 and int(nbt) == 1
    d1d, d2d = nhwc(d1, dtype), nhwc(d2, dtype)
    bs = ops.bn_bwd_reduce(xd, d1d, d2d, scale, shift, mean, invstd, ACT_LRELU, ACT_RELU)
    dx = torch.empty_like(xd)
    dg, db = torch.zeros(c, device=DEV), torch.zeros(c, device=DEV)
    ops.bn_bwd_apply(xd, d1d, d2d, scale, shift, mean, invstd, ACT_LRELU, ACT_RELU, bs, n * h * w, dx, dg, db)
    tolg = 2e-5 if dtype == torch.float32 else 8e-3
    assert rel(from_nhwc(dx), xr.grad) < tolg
    assert rel(dg, gr.grad) < 1e-4 and rel(db, br.grad) < 1e-4


def test_bn_eval_and_dropout_mask():
    g = torch.Generator().manual_seed(9)
    c, n, h, w = 64, 2, 4, 4
    x = torch.randn(n, c, h, w, generator=g)
    gamma, beta = torch.rand(c, generator=g) + 0.5, torch.randn(c, generator=g)
    rm, rv = torch.randn(c, generator=g), torch.rand(c, generator=g) + 0.5
    z = F.batch_norm(x, rm, rv, gamma, beta, training=False, eps=1e-5)
    mask = (torch.rand(n, c, h, w, generator=g) > 0.5)
    ref = F.relu(z * mask * 2.0)
    xd = nhwc(x, torch.float32)
    scale, shift, _, _ = ops.bn_finalize(None, 1.0, gamma.to(DEV), beta.to(DEV), rm.to(DEV), rv.to(DEV), None, 0.1, 1e-5,
                                         c, torch.device(DEV))
    o = torch.empty_like(xd)
    md = mask.permute(0, 2, 3, 1).contiguous().to(DEV, torch.uint8)
    ops.bn_act(xd, scale, shift, ACT_RELU, o, drop_mask=md)
    assert rel(from_nhwc(o), ref) < 1e-5
    # Philox mask: ~half kept, identical when recomputed
    o1, o2 = torch.empty_like(xd), torch.empty_like(xd)
    ops.bn_act(xd, None, None, ACT_NONE, o1, drop_seed=1234)
    ops.bn_act(xd, None, None, ACT_NONE, o2, drop_seed=1234)
    assert torch.equal(o1, o2)
    kept = (o1 != 0).float().mean().item()
    assert 0.4 < kept < 0.6
    assert torch.allclose(o1[o1 != 0], (2 * xd)[o1 != 0])


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("hw", [(8, 8), (9, 7)])
def test_maxpool(dtype, hw):
    g = torch.Generator().manual_seed(1)
    x = torch.randn(2, 64, *hw, generator=g)
    x[0, :, :2, :2] = 0.0  # ties: first max wins
    if dtype == torch.bfloat16:
        x = x.bfloat16().float()
    xr = x.clone().requires_grad_(True)
    y = F.max_pool2d(xr, 2)
    dy = torch.randn(y.shape, generator=g)
    y.backward(dy)
    xd = nhwc(x, dtype)
    yd = ops.maxpool2_fwd(xd)
    assert torch.equal(from_nhwc(yd), y.detach())
    dxd = torch.full_like(xd, 7.0)
    ops.maxpool2_bwd(xd, nhwc(dy, dtype), dxd, accumulate=False)
    assert rel(from_nhwc(dxd), xr.grad) < (1e-6 if dtype == torch.float32 else 5e-3)
    base = torch.ones_like(xd)
    ops.maxpool2_bwd(xd, nhwc(dy, dtype), base, accumulate=True)
    assert rel(from_nhwc(base), xr.grad + 1) < (1e-6 if dtype == torch.float32 else 5e-3)


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float32])
@pytest.mark.parametrize("shape", [(2, 64, 32, 48), (3, 32, 6, 10), (1, 512, 2, 2)])
def test_bn_act_pool_fused(shape, dtype):
    """BatchNorm apply + ReLU + MaxPool2d(2) in one pass (unet_parts.py:21 -> :34) == the two separate kernels, bit for
    bit, with the activation written into a channel slice of a concat buffer."""
    g = torch.Generator().manual_seed(31)
    n, c, h, w = shape
    y = torch.randn(n, h, w, c, generator=g).to(DEV, dtype)
    scale = (torch.rand(c, generator=g) + 0.5).to(DEV)
    shift = (torch.randn(c, generator=g) * 0.3).to(DEV)
    cat_a = torch.full((n, h, w, 2 * c), 7.0, device=DEV, dtype=dtype)
    cat_b = torch.full((n, h, w, 2 * c), 7.0, device=DEV, dtype=dtype)
    ops.bn_act(y, scale, shift, ACT_RELU, cat_a[..., :c])
    pooled_a = ops.maxpool2_fwd(cat_a[..., :c])
    pooled_b = torch.full((n, h // 2, w // 2, c), -3.0, device=DEV, dtype=dtype)
    ops.bn_act_pool(y, scale, shift, cat_b[..., :c], pooled_b)
    assert torch.equal(cat_a, cat_b)
    assert torch.equal(pooled_a, pooled_b)
    ref = F.max_pool2d(torch.relu(y.float() * scale + shift).permute(0, 3, 1, 2), 2).permute(0, 2, 3, 1)
    assert rel(pooled_b.float(), ref) < (1e-6 if dtype == torch.float32 else 5e-3)
    with pytest.raises(RuntimeError):
        ops.bn_act_pool(y[:, :h - 1], scale, shift, cat_b[:, :h - 1, :, :c], pooled_b)


@pytest.mark.parametrize("shape", [(2, 64, 32, 48), (3, 128, 6, 10), (1, 512, 2, 2), (8, 64, 64, 64)])
def test_bn_bwd_stats_pool_fused(shape):
    """MaxPool2d(2) backward folded into the BatchNorm backward statistics pass == gsb_maxpool2_bwd(accumulate) followed by
    the plain statistics pass: the routed gradient dA bit for bit (incl. ties: ReLU zeros, first maximum wins), the sums
    to fp32 summation order."""
    g = torch.Generator().manual_seed(33)
    n, c, h, w = shape
    y = torch.randn(n, h, w, c, generator=g).to(DEV, torch.bfloat16)
    scale = (torch.rand(c, generator=g) + 0.5).to(DEV)
    shift = (torch.randn(c, generator=g) * 0.3).to(DEV)
    mean = (torch.randn(c, generator=g) * 0.1).to(DEV)
    invstd = (torch.rand(c, generator=g) + 0.5).to(DEV)
    a = torch.empty_like(y)
    ops.bn_act(y, scale, shift, ACT_RELU, a)
    dcat = torch.randn(n, h, w, 2 * c, generator=g).to(DEV, torch.bfloat16)
    dpool = torch.randn(n, h // 2, w // 2, c, generator=g).to(DEV, torch.bfloat16)
    d_ref = dcat.clone()
    ops.maxpool2_bwd(a, dpool, d_ref[..., :c], accumulate=True)
    s_ref, dg_ref, db_ref = ops.bn_bwd_stats(y, d_ref[..., :c], None, scale, shift, mean, invstd, ACT_RELU, ACT_NONE)
    d_fused = dcat.clone()
    s_fused, dg, db = ops.bn_bwd_stats(y, d_fused[..., :c], None, scale, shift, mean, invstd, ACT_RELU, ACT_NONE, pool_grad=dpool)
    assert torch.equal(d_ref, d_fused)
    assert rel(s_fused, s_ref) < 2e-6 and rel(dg, dg_ref) < 2e-6 and rel(db, db_ref) < 2e-6
    # against torch: the gradient that reaches the BatchNorm output through relu and max_pool2d
    yt = (y.float() * scale + shift).permute(0, 3, 1, 2).clone().requires_grad_(True)
    at = torch.relu(yt).bfloat16().float()       # argmax on the stored (bf16) activations
    at.retain_grad()
    F.max_pool2d(at, 2).backward(dpool.float().permute(0, 3, 1, 2))
    routed = at.grad.permute(0, 2, 3, 1)
    want = (dcat[..., :c].float() + routed).bfloat16()
    assert torch.equal(d_fused[..., :c], want)


@pytest.mark.parametrize("hw,out_hw", [((4, 4), (8, 8)), ((4, 5), (9, 11)), ((1, 3), (2, 6)), ((16, 16), (32, 32))])
def test_upsample_into_concat_slice(hw, out_hw):
    g = torch.Generator().manual_seed(2)
    n, c = 2, 64
    x = torch.randn(n, c, *hw, generator=g)
    xr = x.clone().requires_grad_(True)
    up = F.interpolate(xr, scale_factor=2, mode="bilinear", align_corners=True)
    dy_, dx_ = out_hw[0] - up.shape[2], out_hw[1] - up.shape[3]
    ref = F.pad(up, [dx_ // 2, dx_ - dx_ // 2, dy_ // 2, dy_ - dy_ // 2])
    dref = torch.randn(ref.shape, generator=g)
    ref.backward(dref)
    xd = nhwc(x, torch.float32)
    cat = torch.full((n, out_hw[0], out_hw[1], 2 * c), 5.0, device=DEV)
    ops.upsample2_fwd(xd, cat[..., c:], dy_ // 2, dx_ // 2)
    assert rel(from_nhwc(cat[..., c:]), ref) < 1e-6
    assert (cat[..., :c] == 5.0).all()
    dcat = torch.zeros_like(cat)
    dcat[..., c:] = nhwc(dref, torch.float32)
    dxd = torch.empty_like(xd)
    ops.upsample2_bwd(dcat[..., c:], dxd, dy_ // 2, dx_ // 2)
    assert rel(from_nhwc(dxd), xr.grad) < 1e-5


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("shape", [(2, 64, 128, 128), (3, 512, 16, 16), (2, 8, 7, 300), (1, 64, 1, 1), (2, 128, 2, 3)])
def test_upsample_block_kernels(shape, dtype):
    """The 2x2-block forward and the separable backward (un-padded x2 case) against torch, at the production shapes
    (unet_parts.py:50: Upsample(scale_factor=2, 'bilinear', align_corners=True)) and at ragged / degenerate ones."""
    g = torch.Generator().manual_seed(21)
    n, c, h, w = shape
    x = torch.randn(n, c, h, w, generator=g)
    if dtype == torch.bfloat16:
        x = x.bfloat16().float()
    xr = x.clone().requires_grad_(True)
    up = F.interpolate(xr, scale_factor=2, mode="bilinear", align_corners=True)
    dref = torch.randn(up.shape, generator=g)
    if dtype == torch.bfloat16:
        dref = dref.bfloat16().float()
    up.backward(dref)
    xd = nhwc(x, dtype)
    out = torch.full((n, 2 * h, 2 * w, c), 5.0, device=DEV, dtype=dtype)
    ops.upsample2_fwd(xd, out, 0, 0)
    tol = 1e-6 if dtype == torch.float32 else 3e-3
    assert rel(from_nhwc(out).float(), up.detach()) < tol
    dxd = torch.full_like(xd, 9.0)
    ops.upsample2_bwd(nhwc(dref, dtype), dxd, 0, 0)
    assert rel(from_nhwc(dxd).float(), xr.grad) < (2e-6 if dtype == torch.float32 else 3e-3)
    # element-wise: no output of the block form is more than a rounding away from torch's
    err = (from_nhwc(out).float() - up.detach()).abs().max().item()
    assert err <= (1e-5 if dtype == torch.float32 else 2.0 ** -8 * up.abs().max().item())


def test_merge_unmerge_upconv():
    from oracle import genseg_oracle as O
    g = torch.Generator().manual_seed(5)
    cin, cout = 6, 5
    st = {f"c._ops._ops.{i}.op.weight": torch.randn(cin, cout, k, k, generator=g) for i, k in enumerate((4, 6, 8))}
    st.update({f"c._ops._ops.{i}.op.bias": torch.randn(cout, generator=g) for i in range(3)})
    leaves = {k: v.clone().requires_grad_(True) for k, v in st.items()}
    alpha = torch.randn(3, generator=g).requires_grad_(True)
    W, b = O.merged_upconv_weight(leaves, "c", alpha, bias=True)
    gW, gb = torch.randn(W.shape, generator=g), torch.randn(b.shape, generator=g)
    ((W * gW).sum() + (b * gb).sum()).backward()
    d = {k: v.to(DEV) for k, v in st.items()}
    ws = [d[f"c._ops._ops.{i}.op.weight"] for i in range(3)]
    bs = [d[f"c._ops._ops.{i}.op.bias"] for i in range(3)]
    ad = alpha.detach().to(DEV)
    wm, bm = ops.merge_upconv(*ws, *bs, ad)
    assert rel(wm, W) < 1e-6 and rel(bm, b) < 1e-6
    dal = torch.zeros(3, device=DEV)
    outs = ops.unmerge_upconv(gW.to(DEV), gb.to(DEV), *ws, *bs, ad, dal)
    for i in range(3):
        assert rel(outs[i], leaves[f"c._ops._ops.{i}.op.weight"].grad) < 1e-6
        assert rel(outs[3 + i], leaves[f"c._ops._ops.{i}.op.bias"].grad) < 1e-6
    assert rel(dal, alpha.grad) < 1e-5


def test_optimizers_match_torch():
    g = torch.Generator().manual_seed(4)
    p0 = torch.randn(10007, generator=g)
    grads = [torch.randn(10007, generator=g) for _ in range(3)]
    pr = p0.clone().requires_grad_(True)
    opt = torch.optim.RMSprop([pr], lr=1e-3, weight_decay=1e-8, momentum=0.9)
    pd = p0.to(DEV)
    sq, mb = torch.zeros_like(pd), torch.zeros_like(pd)
    for gr in grads:
        pr.grad = gr.clone()
        opt.step()
        ops.rmsprop_step(pd, gr.to(DEV), sq, mb, 1e-3, 0.99, 1e-8, 1e-8, 0.9)
    assert rel(pd, pr) < 1e-6
    pr = p0.clone().requires_grad_(True)
    opt = torch.optim.Adam([pr], lr=2e-4, betas=(0.5, 0.999))
    pd = p0.to(DEV)
    m, v = torch.zeros_like(pd), torch.zeros_like(pd)
    for t, gr in enumerate(grads, 1):
        pr.grad = gr.clone()
        opt.step()
        ops.adam_step(pd, gr.to(DEV), m, v, 2e-4, 0.5, 0.999, 1e-8, 0.0, t)
    assert rel(pd, pr) < 1e-6


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("cfg", [(1, 64, 3, 1, 1, 2, 20, 28), (3, 64, 3, 1, 1, 2, 17, 16), (1, 64, 4, 2, 1, 2, 32, 32),
                                 (2, 64, 4, 2, 1, 3, 16, 24)])
def test_stem_conv(dtype, cfg):
    cin, cout, k, s, p, n, h, w = cfg
    g = torch.Generator().manual_seed(11)
    x = torch.rand(n, cin, h, w, generator=g)
    wt = torch.randn(cout, cin, k, k, generator=g) * 0.3
    b = torch.randn(cout, generator=g)
    wr = wt.clone().requires_grad_(True)
    y = F.conv2d(x, wr, b, stride=s, padding=p)
    dy = torch.randn(y.shape, generator=g)
    if dtype == torch.bfloat16:
        dy = dy.bfloat16().float()
    y.backward(dy)
    yd = torch.empty((n, y.shape[2], y.shape[3], cout), device=DEV, dtype=dtype)
    ops.stem_conv_fwd(x.to(DEV), wt.to(DEV), b.to(DEV), yd, k, s, p)
    assert rel(from_nhwc(yd), y) < (1e-5 if dtype == torch.float32 else 4e-3)
    dw = ops.stem_conv_wgrad(x.to(DEV), nhwc(dy, dtype), k, s, p)
    assert rel(dw, wr.grad) < 2e-5


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("c", [1, 2, 3])
def test_head_conv(dtype, c):
    g = torch.Generator().manual_seed(12)
    n, h, w, k = 2, 19, 23, 64
    x = torch.randn(n, k, h, w, generator=g)
    if dtype == torch.bfloat16:
        x = x.bfloat16().float()
    wt = torch.randn(c, k, 1, 1, generator=g) * 0.2
    b = torch.randn(c, generator=g)
    xr, wr, br = x.clone().requires_grad_(True), wt.clone().requires_grad_(True), b.clone().requires_grad_(True)
    y = F.conv2d(xr, wr, br)
    dy = torch.randn(y.shape, generator=g)
    y.backward(dy)
    cat = torch.zeros((n, h, w, 2 * k), device=DEV, dtype=dtype)     # strided (channel-slice) input view
    xd = cat[..., k:]
    xd.copy_(nhwc(x, dtype))
    out = ops.head_conv_fwd(xd, wt.to(DEV), b.to(DEV))
    assert rel(out, y) < 1e-5
    dx, dw, db = ops.head_conv_bwd(xd, wt.to(DEV), dy.to(DEV), True, True, True)
    assert rel(from_nhwc(dx), xr.grad) < (1e-5 if dtype == torch.float32 else 4e-3)
    assert rel(dw, wr.grad) < 2e-5 and rel(db, br.grad) < 1e-5


def test_multi_tensor_optimizers_match_torch():
    """One-launch multi-tensor RMSprop / Adam (the path optim.RMSprop / optim.Adam take) against torch.optim on CPU,
    over tensors whose sizes straddle the 2048-element block granularity."""
    from semantic_segmentation_b200.optim import Adam, RMSprop
    g = torch.Generator().manual_seed(31)
    shapes = [(3,), (2048,), (2049,), (64, 33, 3, 3), (1,), (5000,)]
    p0 = [torch.randn(s, generator=g) for s in shapes]
    grads = [[torch.randn(s, generator=g) for s in shapes] for _ in range(3)]
    for kind in ("rmsprop", "adam"):
        ref = [p.clone().requires_grad_(True) for p in p0]
        dev = [p.clone().to(DEV).requires_grad_(True) for p in p0]
        if kind == "rmsprop":
            ro = torch.optim.RMSprop(ref, lr=1e-3, weight_decay=1e-8, momentum=0.9)
            do = RMSprop(dev, lr=1e-3, weight_decay=1e-8, momentum=0.9)
        else:
            ro = torch.optim.Adam(ref, lr=2e-4, betas=(0.5, 0.999))
            do = Adam(dev, lr=2e-4, betas=(0.5, 0.999))
        for gs in grads:
            for r, d, gr in zip(ref, dev, gs):
                r.grad = gr.clone()
                d.grad = gr.to(DEV)
            ro.step()
            do.step()
        for r, d in zip(ref, dev):
            assert rel(d, r) < 1e-6, kind
        # torch.optim.Optimizer protocol: the state of the fused optimiser loads into torch.optim's class (and a fresh fused
        # one), and both continue identically
        ref2 = [d.detach().cpu().clone().requires_grad_(True) for d in dev]
        dev2 = [d.detach().clone().requires_grad_(True) for d in dev]
        ro2 = type(ro)(ref2, **{k: v for k, v in ro.defaults.items() if k in ("lr", "weight_decay", "momentum", "betas")})
        do2 = type(do)(dev2, lr=9.9)
        ro2.load_state_dict(do.state_dict())
        do2.load_state_dict(copy.deepcopy(do.state_dict()))     # load_state_dict aliases same-device state tensors
        for r, d, d2, gr in zip(ref2, dev, dev2, grads[0]):
            r.grad, d.grad, d2.grad = gr.clone(), gr.to(DEV), gr.to(DEV)
        ro2.step()
        do.step()
        do2.step()
        for r, d, d2 in zip(ref2, dev, dev2):
            assert torch.equal(d.detach(), d2.detach()), kind
            assert rel(d, r) < 1e-6, kind


def test_batched_pack_and_unpack_match_single():
    g = torch.Generator().manual_seed(32)
    ws = [torch.randn(s, generator=g).to(DEV) for s in [(64, 128, 3, 3), (8, 16, 4, 4), (128, 64, 2, 2), (1, 7, 8, 8)]]
    for dtype in (torch.bfloat16, torch.float32):
        jobs, singles = [], []
        for i, w in enumerate(ws):
            rfd = i % 2 == 0
            d0, d1, kh, kw = w.shape
            rows, cols = (d0, d1) if rfd else (d1, d0)
            out = torch.empty((kh * kw, rows, cols), device=DEV, dtype=dtype)
            jobs.append((w, rfd, out))
            singles.append(ops.pack_weight(w, rfd, dtype))
        ops.pack_weight_multi(jobs, dtype)
        for (_, _, out), s in zip(jobs, singles):
            assert torch.equal(out, s)
    ujobs, usingles = [], []
    for i, w in enumerate(ws):
        rfd = i % 2 == 0
        d0, d1, kh, kw = w.shape
        rows, cols = (d0, d1) if rfd else (d1, d0)
        nsl = 1 + i                                # partial-sum slices, added in slice order while un-packing
        dwp = torch.randn((nsl, kh * kw, rows, cols), generator=g).to(DEV)
        out = torch.empty(tuple(w.shape), device=DEV, dtype=torch.float32)
        ujobs.append((dwp, rfd, out, nsl))
        usingles.append(ops.unpack_wgrad(dwp, tuple(w.shape), rfd, nslices=nsl))
        total = dwp.double().sum(0).float()        # slices are added in slice order with fp64 accumulation
        assert rel(usingles[-1], ops.unpack_wgrad(total, tuple(w.shape), rfd)) < 1e-7
        acc = torch.zeros_like(total).reshape(-1)
        ops.sum_slices(dwp, nsl, total.numel(), acc)
        assert rel(acc, dwp.double().sum(0).reshape(-1)) < 1e-6
    ops.unpack_wgrad_multi(ujobs)
    for (_, _, out, _n), s in zip(ujobs, usingles):
        assert torch.equal(out, s)


@pytest.mark.gpu
@pytest.mark.parametrize("shape", [(5, 70, 3, 3), (64, 1, 4, 4), (33, 129, 8, 8), (512, 256, 1, 1), (16, 32, 2, 2), (130, 17, 4, 4),
                                   # full tiles -> the 16-byte vector paths: all taps per tile (9, 16, 1), tap chunks (64 = 4 x 16),
                                   # and mixes of full / ragged tiles and of aligned / unaligned row lengths
                                   (64, 128, 3, 3), (128, 64, 3, 3), (96, 64, 4, 4), (64, 96, 8, 8), (64, 64, 1, 1), (48, 40, 3, 3),
                                   (64, 36, 3, 3), (36, 64, 3, 3), (32, 33, 4, 4), (64, 64, 5, 5)])
def test_pack_unpack_against_torch_permute(shape):
    """Tiled transposes (csrc/layout.cu) against torch.permute: ragged tiles in every dimension, 1..64 taps,
    both row conventions, bf16 and fp32 packs, accumulate mode of the un-pack."""
    g = torch.Generator().manual_seed(sum(shape))
    d0, d1, kh, kw = shape
    w = torch.randn(shape, generator=g).to(DEV)
    for rfd in (True, False):
        ref = w.reshape(d0, d1, kh * kw).permute(2, 0, 1) if rfd else w.reshape(d0, d1, kh * kw).permute(2, 1, 0)
        for dt in (torch.float32, torch.bfloat16):
            assert torch.equal(ops.pack_weight(w, rfd, dt), ref.contiguous().to(dt))
        back = ops.unpack_wgrad(ref.contiguous(), shape, rfd)
        assert torch.equal(back, w)
        acc = torch.ones(shape, device=DEV)
        ops.unpack_wgrad(ref.contiguous(), shape, rfd, out=acc)
        assert torch.equal(acc, w + 1.0)


DET_GEOMS = [("conv", 3, 1, 1, 128, 128, 8, 32, 32, 1), ("conv", 3, 1, 1, 64, 64, 2, 64, 64, 1), ("conv", 4, 2, 1, 64, 128, 5, 4, 4, 1),
             ("convT", 8, 2, 3, 128, 64, 2, 4, 4, 1), ("conv", 3, 1, 1, 64, 128, 3, 9, 11, 0), ("convT", 8, 2, 3, 128, 1, 2, 12, 10, 0),
             ("conv", 4, 2, 1, 2, 64, 2, 16, 12, 0)]


@pytest.mark.parametrize("geom", DET_GEOMS, ids=[f"{g[0]}k{g[1]}s{g[2]}_{g[4]}to{g[5]}_{g[7]}x{g[8]}_tc{g[9]}" for g in DET_GEOMS])
def test_conv_kernels_are_deterministic(geom):
    """No atomics in the split reductions: the weight gradient (tcgen05 halo / small-tile kernels, SIMT, thin), the
    split-K forward of launches with few tiles and the bias gradient give the same BITS on every run."""
    kind, k, stride, pad, cin, cout, n, h, w, use_tc = geom
    dtype = torch.bfloat16 if use_tc else torch.float32
    g = torch.Generator().manual_seed(11)
    x = torch.randn(n, h, w, cin, generator=g).to(DEV, dtype)
    wshape = (cout, cin, k, k) if kind == "conv" else (cin, cout, k, k)
    wt = (torch.randn(wshape, generator=g) / math.sqrt(cin * k * k)).to(DEV)
    fwd_mode = MODE_CONV if kind == "conv" else MODE_CONVT
    oh = (h + 2 * pad - k) // stride + 1 if kind == "conv" else (h - 1) * stride - 2 * pad + k
    ow = (w + 2 * pad - k) // stride + 1 if kind == "conv" else (w - 1) * stride - 2 * pad + k
    dy = torch.randn(n, oh, ow, cout, generator=g).to(DEV, dtype)
    wp = ops.pack_weight(wt, kind == "conv", dtype)
    runs = []
    for _ in range(3):
        y = torch.empty((n, oh, ow, cout), device=DEV, dtype=dtype)
        ops.conv_gather(view_nhwc(x), wp, None, view_nhwc(y), fwd_mode, k, stride, pad, use_tc=use_tc)
        dwp = torch.zeros((k * k, cout, cin), device=DEV, dtype=torch.float32)
        ops.conv_wgrad_into(view_nhwc(x), view_nhwc(dy), dwp, fwd_mode, k, stride, pad, use_tc=use_tc)
        db = torch.empty(cout, device=DEV, dtype=torch.float32)
        ops.bias_grad(view_nhwc(dy), db)
        runs.append((y.clone(), dwp.clone(), db.clone()))
    for r in runs[1:]:
        assert torch.equal(r[0], runs[0][0]) and torch.equal(r[1], runs[0][1]) and torch.equal(r[2], runs[0][2])
    assert rel(runs[0][2], dy.double().sum((0, 1, 2))) < (1e-6 if dtype == torch.float32 else 1e-5)


@pytest.mark.parametrize("shape", [(8, 1, 256, 256), (2, 3, 70, 90), (2, 1, 30, 30)])
def test_bias_grad_planar_views(shape):
    """Bias gradient of the 1..3-channel fp32 NCHW network outputs (generator's outermost ConvTranspose2d, the PatchGAN's last
    convolution: networks.py:590,660): parallel over pixel chunks, partial rows added in order -> deterministic."""
    from semantic_segmentation_b200._lib import view_nchw
    g = torch.Generator().manual_seed(17)
    d = torch.randn(shape, generator=g).to(DEV)
    outs = []
    for _ in range(2):
        db = torch.full((shape[1],), 9.0, device=DEV)
        ops.bias_grad(view_nchw(d), db)
        outs.append(db.clone())
    assert torch.equal(outs[0], outs[1])
    assert rel(outs[0], d.double().sum((0, 2, 3))) < 1e-6


THIN = [  # kind, k, stride, pad, cin, cout, n, h, w : 1..4 channels on one side (thin_conv.cu)
    ("convT", 8, 2, 3, 128, 1, 2, 12, 10),
    ("conv", 4, 1, 1, 512, 1, 2, 9, 9),
    ("conv", 4, 2, 1, 2, 64, 2, 16, 12),
    ("convT", 8, 2, 3, 64, 3, 1, 6, 6),
]


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("geom", THIN, ids=[f"{g[0]}k{g[1]}s{g[2]}_{g[4]}to{g[5]}" for g in THIN])
def test_thin_convolutions(geom, dtype):
    """Layers with 1..4 channels on one side run through thin_gather / thin_scatter / thin_wgrad (use_tc=0 path)."""
    res = conv_case(*geom, dtype=dtype, use_tc=0)
    tol = 1e-5 if dtype == torch.float32 else 6e-3
    assert res["fwd"] < tol and res["dgrad"] < tol and res["wgrad"] < (2e-5 if dtype == torch.float32 else 1e-4), res


@pytest.mark.parametrize("shape", [(2, 128, 12, 10), (1, 64, 9, 17), (8, 128, 128, 128)])
def test_thin_tc_outermost_cell(shape, monkeypatch):
    """The generator's outermost up-cell, ConvTranspose2d(CW -> 1, 8x8 s2 p3) (networks.py:590 with the merged cell of
    operations.py:14-39), on the tensor cores (thin_tc.cu): forward (GEMM over the 64 taps + col2im + bias + tanh), data
    gradient and weight gradient against torch on bf16-representable operands, with the fp32 NCHW one-channel tensors the
    engine uses, and against the SIMT thin kernels."""
    from semantic_segmentation_b200._lib import view_nchw
    n, cw, h, w = shape
    g = torch.Generator().manual_seed(n * 1000 + cw + h)
    x = torch.randn(n, cw, h, w, generator=g).bfloat16().float()
    wt = (torch.randn(cw, 1, 8, 8, generator=g) / math.sqrt(cw * 16)).bfloat16().float()
    bias = torch.randn(1, generator=g) * 0.1
    x.requires_grad_(True)
    wt.requires_grad_(True)
    y = torch.tanh(F.conv_transpose2d(x, wt, bias, stride=2, padding=3))
    dz = torch.randn(y.shape, generator=g).bfloat16().float()          # gradient w.r.t. the pre-activation, bf16-representable
    F.conv_transpose2d(x, wt, bias, stride=2, padding=3).backward(dz)
    xd = nhwc(x.detach(), torch.bfloat16)
    wd = wt.detach().to(DEV)
    wp, wpb = ops.pack_weight(wd, False, torch.bfloat16), ops.pack_weight(wd, True, torch.bfloat16)
    bd = bias.to(DEV)
    dzd = dz.to(DEV).contiguous()
    out = {}
    for flag in ("1", "0"):
        monkeypatch.setenv("GSB_THIN_TC", flag)
        yd = torch.full((n, 1, 2 * h, 2 * w), 7.0, device=DEV)
        ops.conv_gather(view_nhwc(xd), wp, bd, view_nchw(yd), MODE_CONVT, 8, 2, 3, act=ACT_TANH)
        dxd = torch.full_like(xd, 3.0)
        ops.conv_gather(view_nchw(dzd), wpb, None, view_nhwc(dxd), MODE_CONV, 8, 2, 3)
        dwp = torch.zeros((64, 1, cw), device=DEV, dtype=torch.float32)
        ops.conv_wgrad_into(view_nhwc(xd), view_nchw(dzd), dwp, MODE_CONVT, 8, 2, 3)
        dw = ops.unpack_wgrad(dwp, (cw, 1, 8, 8), False)
        torch.cuda.synchronize()
        out[flag] = (yd.cpu(), from_nhwc(dxd), dw.cpu())
    for flag in ("1", "0"):
        yd, dx, dw = out[flag]
        res = dict(fwd=rel(yd, y), dgrad=rel(dx, x.grad), wgrad=rel(dw, wt.grad))
        assert res["fwd"] < 1e-5 and res["dgrad"] < 6e-3 and res["wgrad"] < 1e-4, (flag, res)
    assert rel(out["1"][0], out["0"][0]) < 1e-5 and rel(out["1"][2], out["0"][2]) < 1e-4


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_instance_norm_batched_kernels(dtype):
    """InstanceNorm2d(affine=False) (networks.py:34-35) at batch 8 through the per-image mode of the BatchNorm kernels:
    statistics, normalise + LeakyReLU / ReLU outputs and the backward pass against torch, in ONE launch per stage whatever
    the batch size."""
    from semantic_segmentation_b200._lib import load
    g = torch.Generator().manual_seed(41)
    n, c, h, w = 8, 64, 12, 10
    x = (torch.randn(n, c, h, w, generator=g) * 1.3 + 0.4 * torch.randn(n, c, 1, 1, generator=g))
    if dtype == torch.bfloat16:
        x = x.bfloat16().float()
    xr = x.clone().requires_grad_(True)
    z = F.instance_norm(xr, eps=1e-5)
    a1, a2 = F.leaky_relu(z, 0.2), F.relu(z)
    d1, d2 = torch.randn(a1.shape, generator=g), torch.randn(a1.shape, generator=g)
    if dtype == torch.bfloat16:
        d1, d2 = d1.bfloat16().float(), d2.bfloat16().float()
    (a1 * d1 + a2 * d2).sum().backward()
    xd = nhwc(x, dtype)
    centered = dtype == torch.float32
    l0 = load().gsb_launch_count()
    sums = ops.bn_reduce(xd, per_image=True)
    scale, shift, mean, invstd = ops.bn_finalize(sums, h * w, None, None, None, None, None, 0.1, 1e-5, n * c, torch.device(DEV),
                                                 centered=centered)
    o1, o2 = torch.empty_like(xd), torch.empty_like(xd)
    ce = mean if centered else None
    ops.bn_act(xd, scale, shift, ACT_LRELU, o1, ACT_RELU, o2, center=ce, per_image=True)
    d1d, d2d = nhwc(d1, dtype), nhwc(d2, dtype)
    bs, _, _ = ops.bn_bwd_stats(xd, d1d, d2d, scale, shift, mean, invstd, ACT_LRELU, ACT_RELU, 0, None, False, None, center=ce,
                                per_image=True)
    dx = torch.empty_like(xd)
    ops.bn_bwd_apply(xd, d1d, d2d, scale, shift, mean, invstd, ACT_LRELU, ACT_RELU, bs, h * w, dx, center=ce, per_image=True)
    launches = load().gsb_launch_count() - l0
    assert launches == 7          # reduce (2) + finalize + act + bwd partial + bwd stats + bwd apply: independent of the batch size
    tol = 2e-5 if dtype == torch.float32 else 8e-3
    assert rel(from_nhwc(o1), a1) < tol and rel(from_nhwc(o2), a2) < tol
    assert rel(mean.view(n, c), x.mean((2, 3))) < 1e-5
    assert rel(from_nhwc(dx), xr.grad) < (5e-5 if dtype == torch.float32 else 1e-2)


@pytest.mark.parametrize("cfg", [(3, 1, 1, 5, 40, 56), (4, 2, 1, 3, 64, 32), (3, 1, 1, 2, 256, 256)])
def test_stem_conv_tensor_core(cfg):
    """Network-input convolutions on the tensor cores (stem_tc.cu: im2col tile in shared memory, hi / lo bf16 split of the
    fp32 input and weights, tcgen05.mma, fp32 accumulation in TMEM): forward against torch's fp32 convolution to bf16
    output rounding, the fused BatchNorm statistics of the stored values exactly, the weight gradient to 2e-5; a pixel
    count that is not a multiple of the 128-pixel tile; bit-reproducible."""
    k, s, p, n, h, w = cfg
    g = torch.Generator().manual_seed(13)
    x = torch.rand(n, 1, h, w, generator=g)
    wt = torch.randn(64, 1, k, k, generator=g) * 0.3
    wr = wt.clone().requires_grad_(True)
    y = F.conv2d(x, wr, None, stride=s, padding=p)
    dy = torch.randn(y.shape, generator=g).bfloat16().float()
    y.backward(dy)
    oh, ow = y.shape[2], y.shape[3]
    xd, wd = x.to(DEV), wt.to(DEV)
    outs = []
    for _ in range(2):
        yd = torch.empty((n, oh, ow, 64), device=DEV, dtype=torch.bfloat16)
        stats, rows = ops.stem_conv_fwd(xd, wd, None, yd, k, s, p, want_stats=True)
        dw = ops.stem_conv_wgrad(xd, nhwc(dy, torch.bfloat16), k, s, p)
        sums = ops.bn_reduce_partials(stats, rows, 64)
        outs.append((yd.clone(), sums.clone(), dw.clone()))
    yd, sums, dw = outs[0]
    assert all(torch.equal(a, b) for a, b in zip(outs[0], outs[1]))
    assert rel(from_nhwc(yd), y) < 3e-3
    # against the exactly rounded fp32 result: the split keeps 16 mantissa bits of every product -> at most a few ulps of bf16
    exact = y.detach().bfloat16().float()
    assert float((from_nhwc(yd) - exact).abs().max()) <= 2.0 ** -7 * float(y.detach().abs().max())
    yq = yd.double()
    assert rel(sums, torch.cat([yq.sum((0, 1, 2)), (yq * yq).sum((0, 1, 2))])) < 1e-5
    assert rel(dw, wr.grad) < 2e-5
