"""HBM rooflines of the non-convolution kernels of the U-Net step ("HBM GB/s of the non-conv kernels", target
>= 70 % of the measured copy bandwidth).  Every kernel class is timed ALONE with CUDA events on the launching stream, on the
production shape that carries most of its time in the step (the 64-channel 256x256 level unless noted), and its
ALGORITHMIC bytes -- every input element read once, every output element written once, at the storage dtype -- are divided
by the measured duration.  Working sets are far larger than the 126 MB L2 (a [64,256,256,64] bf16 tensor is 537 MB).
Used by bench.py (`roofline_hbm` block of the JSON line); not part of the product path.
"""
import torch


def _time(fn, iters=5, warm=2):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(iters):
        fn()
    e.record()
    torch.cuda.synchronize()
    return s.elapsed_time(e) / iters * 1e-3


def hbm_rooflines(b: int, hbm_gbs: float, n_params: int = 17262976):
    """b = images per GPU.  Returns {kernel: {"bytes", "us", "gbs", "frac", "shape"}}."""
    from semantic_segmentation_b200 import ops
    from semantic_segmentation_b200._lib import ACT_RELU
    dev = torch.device("cuda")
    bf = torch.bfloat16
    H = W = 256
    C = 64
    T = b * H * W * C * 2                      # bytes of one [b,256,256,64] bf16 tensor
    out = {}

    def rec(name, nbytes, t, shape):
        out[name] = {"bytes": int(nbytes), "us": t * 1e6, "gbs": nbytes / t / 1e9, "frac": nbytes / t / 1e9 / hbm_gbs, "shape": shape}

    y = torch.randn(b, H, W, C, device=dev).to(bf)
    dA = torch.randn(b, H, W, C, device=dev).to(bf)
    o = torch.empty_like(y)
    scale = torch.rand(C, device=dev) + 0.5
    shift = torch.randn(C, device=dev) * 0.1
    mean = torch.zeros(C, device=dev)
    invstd = torch.ones(C, device=dev)
    shp = "[%d,256,256,64] bf16" % b
    rec("bn_act (BN apply + ReLU)", 2 * T, _time(lambda: ops.bn_act(y, scale, shift, ACT_RELU, o)), shp)
    rec("bn_bwd_stats (sum dy, sum dy*xhat)", 2 * T,
        _time(lambda: ops.bn_bwd_stats(y, dA, None, scale, shift, mean, invstd, ACT_RELU, 0)), shp)
    sums = torch.zeros(2 * C, device=dev, dtype=torch.float64)
    rec("bn_bwd_apply (dx of BN + ReLU)", 3 * T,
        _time(lambda: ops.bn_bwd_apply(y, dA, None, scale, shift, mean, invstd, ACT_RELU, 0, sums, float(b * H * W), o)), shp)
    rec("bn_reduce (sum x, sum x^2; the conv epilogue does this in the tcgen05 path)", T, _time(lambda: ops.bn_reduce(y)), shp)
    # pooling / upsampling at the first level
    pooled = ops.maxpool2_fwd(y)
    rec("maxpool2_fwd", T + T // 4, _time(lambda: ops.maxpool2_fwd(y)), shp)
    dpool = torch.randn_like(pooled)
    dx = torch.zeros_like(y)
    rec("maxpool2_bwd (accumulating into the skip gradient)", T + T // 4 + 2 * T,
        _time(lambda: ops.maxpool2_bwd(y, dpool, dx, True)), shp)
    cat = torch.empty(b, H, W, 2 * C, device=dev, dtype=bf)
    low = torch.randn(b, H // 2, W // 2, C, device=dev).to(bf)
    rec("upsample2_fwd (bilinear x2 into the concat buffer)", T // 4 + T, _time(lambda: ops.upsample2_fwd(low, cat[..., C:], 0, 0)),
        "[%d,128,128,64] -> [%d,256,256,128][..., 64:]" % (b, b))
    dlow = torch.empty_like(low)
    rec("upsample2_bwd", T + T // 4, _time(lambda: ops.upsample2_bwd(cat[..., C:], dlow, 0, 0)), "same, backward")
    del cat, low, dlow, pooled, dpool, dx
    # network edges
    x = torch.rand(b, 1, H, W, device=dev)
    w0 = torch.randn(C, 1, 3, 3, device=dev) * 0.3
    px = b * H * W
    rec("stem_conv_fwd (1 -> 64, fp32 NCHW in, bf16 NHWC out, fused BN statistics)", px * (4 + C * 2),
        _time(lambda: ops.stem_conv_fwd(x, w0, None, o, 3, 1, 1, want_stats=True)), "[%d,1,256,256] fp32 -> %s" % (b, shp))
    rec("stem_conv_wgrad", px * (4 + C * 2), _time(lambda: ops.stem_conv_wgrad(x, dA, 3, 1, 1)), "same, weight gradient")
    wh = torch.randn(2, C, 1, 1, device=dev) * 0.1
    bh = torch.zeros(2, device=dev)
    logits = ops.head_conv_fwd(y, wh, bh)
    rec("head_conv_fwd (64 -> 2 logits, fp32 NCHW out)", px * (C * 2 + 8), _time(lambda: ops.head_conv_fwd(y, wh, bh)), shp)
    dl = torch.randn_like(logits)
    rec("head_conv_bwd (dx + dw + db in one pass)", px * (2 * C * 2 + 8),
        _time(lambda: ops.head_conv_bwd(y, wh, dl, True, True, True)), shp)
    labels = torch.randint(0, 2, (b, H, W), device=dev)
    rec("seg_loss_reduce (CE + Dice sums)", px * 16, _time(lambda: ops.seg_loss_reduce(logits, labels)), "[%d,2,256,256] fp32 + int64 labels" % b)
    ssum = ops.seg_loss_reduce(logits, labels)
    gs = torch.ones((), device=dev)
    rec("seg_loss_bwd", px * 24, _time(lambda: ops.seg_loss_bwd(logits, labels, ssum, float(px), gs)), "same")
    del x, logits, dl, labels, y, dA, o
    # weight side (batch independent): RMSprop over the 17.26 M parameters
    p = torch.randn(n_params, device=dev)
    g = torch.randn(n_params, device=dev)
    sq = torch.zeros(n_params, device=dev)
    mb = torch.zeros(n_params, device=dev)
    rec("rmsprop (momentum; 4 reads + 3 writes of 4 B per parameter)", n_params * 28,
        _time(lambda: ops.rmsprop_step_multi([p], [g], [sq], [mb], 1e-4, 0.99, 1e-8, 1e-8, 0.9, 1.0)), "%d parameters" % n_params)
    return out
