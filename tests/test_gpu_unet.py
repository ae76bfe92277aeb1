"""GPU parity of the U-Net path (through the reference-shaped API) against the golden fixtures generated
by the reference itself and against the CPU oracle run live on the same seeded inputs."""
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

import semantic_segmentation_b200 as S  # noqa: E402
from oracle import genseg_oracle as O  # noqa: E402
from semantic_segmentation_b200 import ops  # noqa: E402
from semantic_segmentation_b200.unet import UNet  # noqa: E402
from semantic_segmentation_b200.unet.evaluate import evaluate  # noqa: E402
from semantic_segmentation_b200.util.dice_score import dice_coeff, dice_loss, multiclass_dice_coeff, seg_loss  # noqa: E402

DEV = "cuda"
CASES = [
    ("unet_c2_bilinear_32", 1, 2, True, 2, 32, 32, 1, 2),
    ("unet_c1_convT_32", 1, 1, False, 2, 32, 32, 3, 4),
    ("unet_c2_bilinear_odd", 1, 2, True, 2, 36, 44, 5, 6),
    ("unet_c3_rgb_convT_48", 3, 3, False, 3, 48, 32, 7, 8),
]


def rel(a, b):
    a = torch.as_tensor(a).detach().double().cpu().flatten()
    b = torch.as_tensor(b).detach().double().cpu().flatten()
    return float((a - b).norm() / b.norm().clamp_min(1e-30))


def run_case(case, precision, golden_dir):
    name, nch, ncls, bil, n, h, w, wseed, xseed = case
    gold = np.load(os.path.join(golden_dir, name + ".npz"))
    S.set_precision(precision)
    try:
        net = UNet(nch, ncls, bil)
        net.load_state_dict(O.seeded_state(O.unet_state_spec(nch, ncls, bil), wseed), strict=True)
        net = net.to(DEV).train()
        x, y = O.seeded_inputs(xseed, n, nch, h, w, ncls)
        xd = x.to(DEV).requires_grad_(True)
        logits = net(xd)
        logits.retain_grad()
        total, ce, dl = seg_loss(logits, y.to(DEV), return_parts=True)
        total.backward()
        torch.cuda.synchronize()
        out = dict(logits=rel(logits, gold["logits"]), loss=abs(total.item() - float(gold["loss"])),
                   ce=abs(ce.item() - float(gold["ce"])), dice=abs(dl.item() - float(gold["dice"])),
                   dlogits=rel(logits.grad, gold["dlogits"]), dx=rel(xd.grad, gold["dx"]))
        params = dict(net.named_parameters())
        keys = [str(k) for k in gold["grad_keys"]]
        norms = np.array([params[k].grad.norm().item() for k in keys])
        out["grad_norms"] = float(np.max(np.abs(norms - gold["grad_norms"]) / (gold["grad_norms"] + 1e-12)))
        worst = 0.0
        for f in gold.files:
            if f.startswith("grad:"):
                worst = max(worst, rel(params[f[5:]].grad, gold[f]))
        out["grads"] = worst
        sd = net.state_dict()
        worst = 0.0
        for f in gold.files:
            if f.startswith("after:") and "num_batches" not in f:
                worst = max(worst, rel(sd[f[6:]], gold[f]))
        out["running"] = worst
        assert int(sd["inc.double_conv.1.num_batches_tracked"]) == 1
        net.eval()
        with torch.no_grad():
            le = net(x.to(DEV))
        out["logits_eval"] = rel(le, gold["logits_eval"])
        return out, net, x, y, gold
    finally:
        S.set_precision("bf16")


@pytest.mark.parametrize("case", CASES, ids=[c[0] for c in CASES])
def test_unet_fp32_validation_mode_matches_reference(case, golden_dir):
    """fp32 validation mode: rel-L2 <= 1e-5 on activations / logits / loss / dlogits (north_star, with 2e-5
    head-room for the 3-channel case).  Whole-network parameter gradients get 5e-2: a pre-activation within
    ~1e-6 of zero may take the other ReLU branch than the CPU reference (the forward is unaffected, that
    element's gradient flips; ~0.4 such elements are expected per layer of these fixtures), which moves every
    upstream gradient tensor by 1e-3..1e-2 rel-L2; the layer-local tests in test_gpu_ops.py hold 1e-5 on every
    kernel's gradients, and tools/grad_diag.py shows 1e-5 on whole-network gradients when no flip occurs."""
    out, *_ = run_case(case, "fp32", golden_dir)
    print(case[0], "fp32", out)
    assert out["logits"] < 2e-5 and out["logits_eval"] < 1e-5, out
    assert out["loss"] < 1e-5 and out["ce"] < 1e-5 and out["dice"] < 1e-5, out
    assert out["dlogits"] < 1e-5 and out["running"] < 1e-5, out
    # measured with tools/unet_diag.py: errors are ~1e-5 up to the first layer where ONE pre-activation changes
    # ReLU branch, where d(beta) jumps (e.g. 4e-4 / 1.6e-2) and everything upstream inherits it
    assert out["dx"] < 5e-2 and out["grads"] < 5e-2 and out["grad_norms"] < 5e-2, out


@pytest.mark.parametrize("case", CASES, ids=[c[0] for c in CASES])
def test_unet_bf16_matches_reference(case, golden_dir):
    """bf16 production mode (tcgen05 convs), whole network vs the fp32 reference: loss within 5e-3, logits
    within 1e-1 (18 conv+BN layers of bf16 storage drift ~0.3-0.5 % each; measured 4-7e-2 on these random-weight
    fixtures), eval-mode logits within 3e-2, dlogits within 5e-2.  The <= 1e-2 per-layer bound of the north_star
    is checked layer-locally (identical inputs) in test_gpu_ops.py / test_double_conv_bf16_layer_local: whole-
    network gradients are dominated by ReLU-branch flips caused by the forward drift."""
    out, *_ = run_case(case, "bf16", golden_dir)
    print(case[0], "bf16", out)
    assert out["logits"] < 1e-1 and out["logits_eval"] < 3e-2, out
    assert out["loss"] < 5e-3, out
    assert out["dlogits"] < 5e-2 and out["running"] < 3e-2, out


def test_seg_loss_and_dice_api_match_oracle(golden_dir):
    g = np.load(os.path.join(golden_dir, "dice_kats.npz"))
    dev = DEV
    a = torch.tensor([[1., 1.], [0., 0.]], device=dev)
    b = torch.tensor([[1., 0.], [1., 0.]], device=dev)
    assert dice_coeff(torch.ones(1, 2, 2, device=dev), torch.ones(1, 2, 2, device=dev)).item() == 1.0
    assert dice_coeff(torch.zeros(1, 2, 2, device=dev), torch.zeros(1, 2, 2, device=dev)).item() == 1.0
    assert dice_coeff(torch.ones(1, 2, 2, device=dev), torch.zeros(1, 2, 2, device=dev)).item() == \
        pytest.approx(float(g["ones_zeros"]), rel=1e-5)
    assert dice_coeff(a[None], b[None]).item() == pytest.approx(float(g["ab"]), abs=1e-7)
    assert dice_loss(torch.stack([a, b]), torch.stack([b, b])).item() == pytest.approx(0.25, abs=1e-7)
    p, t = torch.from_numpy(g["rand_p"]).to(dev), torch.from_numpy(g["rand_t"]).to(dev)
    assert multiclass_dice_coeff(p, t, False).item() == pytest.approx(float(g["rand_mc_false"]), abs=1e-6)
    assert multiclass_dice_coeff(p, t, True).item() == pytest.approx(float(g["rand_mc_true"]), abs=1e-6)
    assert dice_loss(p, t, True).item() == pytest.approx(float(g["rand_loss_mc"]), abs=1e-6)
    with pytest.raises(AssertionError):
        dice_loss(torch.ones(4, 4, device=dev), torch.ones(4, 4, device=dev))
    # gradient of dice_loss vs oracle autograd
    pr = p.clone().requires_grad_(True)
    dice_loss(pr, t, True).backward()
    pc = p.cpu().clone().requires_grad_(True)
    O.dice_loss(pc, t.cpu(), True).backward()
    assert rel(pr.grad, pc.grad) < 1e-5
    # fused seg loss vs oracle composition, C = 3 and C = 1
    for nm in ("ev3", "ev1"):
        lg = torch.from_numpy(g[nm + "_logits"])
        y = torch.from_numpy(g[nm + "_y"])
        lc = lg.clone().requires_grad_(True)
        tot, ce, dl = O.seg_loss(lc, y)
        tot.backward()
        ld = lg.to(dev).requires_grad_(True)
        tot_d, ce_d, dl_d = seg_loss(ld, y.to(dev), return_parts=True)
        (tot_d * 1.0).backward()
        assert abs(tot_d.item() - tot.item()) < 1e-5 and abs(ce_d.item() - ce.item()) < 1e-5
        assert abs(dl_d.item() - dl.item()) < 1e-5
        assert rel(ld.grad, lc.grad) < 1e-5, nm


def test_eval_counts_bit_exact(golden_dir):
    """argmax masks and Dice intersection/union counts from identical logits: bit-exact (north_star)."""
    g = np.load(os.path.join(golden_dir, "dice_kats.npz"))
    for nm in ("ev3", "ev1"):
        lg, y = g[nm + "_logits"], g[nm + "_y"]
        ref = O.eval_counts(lg, y)
        n, c = lg.shape[0], lg.shape[1]
        counts, mask = ops.eval_counts(torch.from_numpy(lg).to(DEV), torch.from_numpy(y).to(DEV).reshape(n, *lg.shape[2:]))
        counts = counts.cpu().numpy()
        assert np.array_equal(mask.cpu().numpy(), ref["mask"])
        assert np.array_equal(counts[..., 0], ref["inter"]) and np.array_equal(counts[..., 1], ref["psum"])
        assert np.array_equal(counts[..., 2], ref["tsum"])
    # larger random case incl. NaN and ties
    gen = torch.Generator().manual_seed(0)
    lg = torch.randn(3, 4, 64, 48, generator=gen)
    lg[1, :, 5, :7] = 1.0
    lg[2, 2, 0, 0] = float("nan")
    y = torch.randint(0, 4, (3, 64, 48), generator=gen)
    ref = O.eval_counts(lg.numpy(), y.numpy())
    counts, mask = ops.eval_counts(lg.to(DEV), y.to(DEV))
    assert np.array_equal(mask.cpu().numpy(), ref["mask"])
    assert np.array_equal(counts.cpu().numpy()[..., 0], ref["inter"])


@pytest.mark.parametrize("case", CASES[:2], ids=[c[0] for c in CASES[:2]])
def test_evaluate_matches_reference(case, golden_dir):
    name, nch, ncls, bil, n, h, w, wseed, xseed = case
    out, net, x, y, gold = run_case(case, "fp32", golden_dir)
    S.set_precision("fp32")
    try:
        with torch.no_grad():
            d = evaluate(net, [dict(image=x, mask=y)], torch.device(DEV), False)
            d_sharded = evaluate(net, [dict(image=x, mask=y)], torch.device(DEV), False, sharded=True)
        assert float(d) == float(d_sharded)   # one process: the sharded form scores every batch itself
        assert net.training          # evaluate() leaves the net in train mode like the reference (:42)
        # masks from OUR logits vs oracle counts on OUR logits: bit exact; Dice vs reference's: close
        net.eval()
        with torch.no_grad():
            lg = net(x.to(DEV))
        ref = O.eval_counts(lg.cpu().numpy(), y.numpy())
        counts, _ = ops.eval_counts(lg, y.to(DEV).reshape(n, h, w))
        assert np.array_equal(counts.cpu().numpy()[..., 0], ref["inter"])
        assert abs(float(d) - float(gold["eval_dice"])) < 2e-3
    finally:
        S.set_precision("bf16")


def test_autograd_contract():
    """retain_graph, autograd.grad(allow_unused), requires_grad_(False) skipping wgrad, no_grad."""
    S.set_precision("fp32")
    try:
        net = UNet(1, 2, True).to(DEV)
        x = torch.rand(2, 1, 32, 32, device=DEV)
        y = torch.randint(0, 2, (2, 32, 32), device=DEV)
        loss = seg_loss(net(x), y)
        ps = [p for p in net.parameters()]
        g1 = torch.autograd.grad(loss, ps, retain_graph=True, allow_unused=True)
        g2 = torch.autograd.grad(loss, ps, allow_unused=True)
        for a, b in zip(g1, g2):
            assert a is not None and rel(a, b) < 1e-5
        for p in net.inc.parameters():
            p.requires_grad_(False)
        loss = seg_loss(net(x), y)
        loss.backward()
        assert all(p.grad is None for p in net.inc.parameters())
        assert net.outc.conv.weight.grad is not None
        with torch.no_grad():
            out = net(x)
        assert not out.requires_grad
        with pytest.raises(RuntimeError):
            net.cpu()(x.cpu())
    finally:
        S.set_precision("bf16")


def test_standalone_parts_match_torch():
    """DoubleConv / Down / Up / OutConv used on their own (same names + state_dict as the reference)."""
    import torch.nn.functional as F
    from semantic_segmentation_b200.unet.unet_parts import DoubleConv, Down, OutConv, Up
    S.set_precision("fp32")
    try:
        torch.manual_seed(0)
        x = torch.randn(2, 64, 16, 16)
        for mod, ref_fn in ((DoubleConv(64, 128), None), (Down(64, 128), None), (OutConv(64, 3), None)):
            mod = mod.train()
            st = {k: v.clone() for k, v in mod.state_dict().items()}
            xc = x.clone().requires_grad_(True)
            if isinstance(mod, OutConv):
                yref = F.conv2d(xc, st["conv.weight"], st["conv.bias"])
            elif isinstance(mod, Down):
                st2 = {"m" + k[len("maxpool_conv.1"):]: v for k, v in st.items()}
                yref = O.double_conv(F.max_pool2d(xc, 2), st2, "m", True, False)
            else:
                st2 = {"m." + k: v for k, v in st.items()}
                yref = O.double_conv(xc, st2, "m", True, False)
            yref.square().sum().backward()
            md = mod.to(DEV)
            xd = x.to(DEV).requires_grad_(True)
            yd = md(xd)
            yd.square().sum().backward()
            assert rel(yd, yref) < 1e-5 and rel(xd.grad, xc.grad) < 1e-4, type(mod)
        for bil in (True, False):
            up = Up(128, 64, bil).train()
            st = {"u." + k: v.clone() for k, v in up.state_dict().items()}
            x1, x2 = torch.randn(2, 64 if bil else 128, 8, 8), torch.randn(2, 64, 17, 16)
            a, b = x1.clone().requires_grad_(True), x2.clone().requires_grad_(True)
            yref = O.up_block(a, b, st, "u", bil, True, False)
            yref.square().sum().backward()
            ud = up.to(DEV)
            ad, bd = x1.to(DEV).requires_grad_(True), x2.to(DEV).requires_grad_(True)
            yd = ud(ad, bd)
            yd.square().sum().backward()
            assert rel(yd, yref) < 1e-5 and rel(ad.grad, a.grad) < 1e-4 and rel(bd.grad, b.grad) < 1e-4, bil
    finally:
        S.set_precision("bf16")


def test_double_conv_bf16_layer_local():
    """north_star per-layer bound in bf16: a DoubleConv fed the SAME (bf16-representable) input and weights as
    the fp32 oracle; activations rel-L2 <= 1e-2, parameter / input gradients <= 5e-2 (one BN+ReLU stage of
    branch flips), on the tcgen05 path (64 -> 128 channels)."""
    from semantic_segmentation_b200.unet.unet_parts import DoubleConv
    torch.manual_seed(1)
    mod = DoubleConv(64, 128).train()
    with torch.no_grad():
        for p in mod.parameters():
            p.copy_(p.bfloat16().float())
    st = {"m." + k: v.clone() for k, v in mod.state_dict().items()}
    x = torch.randn(4, 64, 32, 32).bfloat16().float()
    xc = x.clone().requires_grad_(True)
    leaves = {k: v.clone().requires_grad_(True) for k, v in st.items() if v.dtype.is_floating_point and "running" not in k}
    st.update(leaves)
    yref = O.double_conv(xc, st, "m", True, False)
    g = torch.randn(yref.shape, generator=torch.Generator().manual_seed(2)).bfloat16().float()
    yref.backward(g)
    md = mod.to(DEV)
    xd = x.to(DEV).requires_grad_(True)
    yd = md(xd)
    yd.backward(g.to(DEV))
    assert rel(yd, yref) < 1e-2
    assert rel(xd.grad, xc.grad) < 5e-2
    for k, p in md.named_parameters():
        assert rel(p.grad, leaves["m." + k].grad) < 5e-2, k


@pytest.mark.parametrize("shape", [(4, 64, 64), (8, 256, 256)], ids=["4x64", "8x256"])
def test_training_step_is_bit_reproducible(shape):
    """No floating-point atomics anywhere on the path (split weight gradients and split-K convolutions store per-split
    slices that are added in slice order, edge / loss reductions go through per-CTA rows): two runs of the same bf16
    training step -- logits, loss, EVERY parameter gradient, the BatchNorm running statistics and the weights after an
    RMSprop step -- are bit-identical.  64x64 exercises the small-tile kernels (4x4 .. 8x8 levels, split-K), 8x256x256 is
    the per-GPU shape of the 8-GPU bench."""
    from semantic_segmentation_b200.optim import RMSprop
    n, h, w = shape
    g = torch.Generator().manual_seed(5)
    x = torch.rand(n, 1, h, w, generator=g).to(DEV)
    y = torch.randint(0, 2, (n, h, w), generator=g).to(DEV)
    runs = []
    for _ in range(2):
        torch.manual_seed(1234)
        net = UNet(1, 2, True).to(DEV).train()
        opt = RMSprop(net.parameters(), lr=1e-4, weight_decay=1e-8, momentum=0.9)
        logits = net(x)
        loss = seg_loss(logits, y)
        loss.backward()
        grads = [p.grad.clone() for p in net.parameters()]
        opt.step()
        torch.cuda.synchronize()
        runs.append((logits.detach().clone(), loss.detach().clone(), grads, [p.detach().clone() for p in net.parameters()],
                     [b.clone() for b in net.buffers()]))
    a, b = runs
    assert torch.equal(a[0], b[0]) and torch.equal(a[1], b[1])
    for k, (ga, gb) in zip([k for k, _ in UNet(1, 2, True).named_parameters()], zip(a[2], b[2])):
        assert torch.equal(ga, gb), k
    assert all(torch.equal(pa, pb) for pa, pb in zip(a[3], b[3]))
    assert all(torch.equal(pa, pb) for pa, pb in zip(a[4], b[4]))


def test_pool_fusion_matches_separate_kernels(monkeypatch):
    """BatchNorm + ReLU + MaxPool (forward) and MaxPool backward + BatchNorm statistics (backward) fused at the encoder seams
    (GSB_POOL_FUSION, default on) against the separate kernels: logits bit-identical, gradients to summation order."""
    import semantic_segmentation_b200 as S
    from semantic_segmentation_b200.unet import UNet
    from semantic_segmentation_b200.util.dice_score import seg_loss
    S.set_precision("bf16")
    g = torch.Generator().manual_seed(9)
    x = torch.rand(4, 1, 64, 64, generator=g).cuda()
    yl = torch.randint(0, 2, (4, 64, 64), generator=g).cuda()
    out = {}
    for flag in ("1", "0"):
        monkeypatch.setenv("GSB_POOL_FUSION", flag)
        torch.manual_seed(77)
        net = UNet(1, 2, True).cuda().train()
        logits = net(x)
        seg_loss(logits, yl).backward()
        out[flag] = (logits.detach().clone(), {k: p.grad.clone() for k, p in net.named_parameters()})
    assert torch.equal(out["1"][0], out["0"][0])
    for k, gr in out["1"][1].items():
        ref = out["0"][1][k]
        err = float((gr.double() - ref.double()).norm() / ref.double().norm().clamp_min(1e-30))
        assert err < 2e-2, (k, err)     # bf16 step: the statistics differ in fp32 summation order only (measured ~1e-3)


def test_eval_forward_folds_batchnorm_into_the_convolution(monkeypatch):
    """Inference (net.eval(), no_grad): Conv -> BatchNorm(running statistics) -> ReLU runs as ONE kernel (per-channel affine
    epilogue of the tcgen05 kernel, GSB_EVAL_FOLD) -- against the separate kernels: logits within the bf16 rounding of one
    stored intermediate, identical argmax masks up to ties, far fewer launches."""
    import semantic_segmentation_b200 as S
    from semantic_segmentation_b200._lib import load
    from semantic_segmentation_b200.unet import UNet
    S.set_precision("bf16")
    torch.manual_seed(31)
    net = UNet(1, 2, True).cuda()
    g = torch.Generator().manual_seed(8)
    with torch.no_grad():                       # non-trivial running statistics
        for m in net.modules():
            if isinstance(m, torch.nn.BatchNorm2d):
                m.running_mean.copy_(torch.randn(m.num_features, generator=g).cuda() * 0.1)
                m.running_var.copy_((torch.rand(m.num_features, generator=g) + 0.5).cuda())
    net.eval()
    x = torch.rand(4, 1, 128, 128, generator=g).cuda()
    out, launches = {}, {}
    for flag in ("1", "0"):
        monkeypatch.setenv("GSB_EVAL_FOLD", flag)
        with torch.no_grad():
            net(x)
            l0 = load().gsb_launch_count()
            out[flag] = net(x).float()
            launches[flag] = load().gsb_launch_count() - l0
    err = float((out["1"] - out["0"]).norm() / out["0"].norm())
    assert err < 1e-2, err
    agree = (out["1"].argmax(1) == out["0"].argmax(1)).float().mean().item()
    assert agree > 0.995, agree
    assert launches["1"] <= launches["0"] - 8, launches       # the BatchNorm passes of the >= 16 x 16 levels are gone (62 -> 53)
    # and against the fp32 validation mode (separate kernels, centred BatchNorm): the production-dtype bound
    S.set_precision("fp32")
    try:
        with torch.no_grad():
            ref = net(x).float()
    finally:
        S.set_precision("bf16")
    assert float((out["1"] - ref).norm() / ref.norm()) < 6e-2
