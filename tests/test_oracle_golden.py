"""CPU: pin oracle/genseg_oracle.py against the fixtures produced by the reference
itself (tests/golden/make_golden.py).  fp32 on both sides -> tight tolerances."""
import os

import numpy as np
import pytest
import torch
import torch.nn.functional as F

from oracle import genseg_oracle as O

torch.set_num_threads(8)

UNET_CASES = [
    ("unet_c2_bilinear_32", 1, 2, True, 2, 32, 32, 1, 2),
    ("unet_c1_convT_32", 1, 1, False, 2, 32, 32, 3, 4),
    ("unet_c2_bilinear_odd", 1, 2, True, 2, 36, 44, 5, 6),
    ("unet_c3_rgb_convT_48", 3, 3, False, 3, 48, 32, 7, 8),
]


def rel_l2(a, b):
    a = torch.as_tensor(a).double().flatten()
    b = torch.as_tensor(b).double().flatten()
    return float((a - b).norm() / b.norm().clamp_min(1e-30))


@pytest.mark.parametrize("case", UNET_CASES, ids=[c[0] for c in UNET_CASES])
def test_unet_oracle_matches_reference(case, golden_dir):
    name, nch, ncls, bil, n, h, w, wseed, xseed = case
    gold = np.load(os.path.join(golden_dir, name + ".npz"))
    st = O.seeded_state(O.unet_state_spec(nch, ncls, bil), wseed)
    x, y = O.seeded_inputs(xseed, n, nch, h, w, ncls)
    assert abs(x.double().sum().item() - float(gold["x_sum"])) < 1e-6   # RNG drift guard
    assert int(y.sum()) == int(gold["y_sum"])
    params = {k: v.clone().requires_grad_(True) for k, v in st.items() if v.dtype.is_floating_point
              and "running" not in k}
    work = dict(st)
    work.update(params)
    x = x.requires_grad_(True)
    logits = O.unet_forward(work, x, bil, training=True, update_running=True)
    logits.retain_grad()
    loss, ce, dl = O.seg_loss(logits, y)
    loss.backward()
    assert rel_l2(logits.detach(), gold["logits"]) < 1e-5
    assert abs(loss.item() - float(gold["loss"])) < 1e-5
    assert abs(ce.item() - float(gold["ce"])) < 1e-5
    assert abs(dl.item() - float(gold["dice"])) < 1e-5
    assert rel_l2(logits.grad, gold["dlogits"]) < 1e-5
    assert rel_l2(x.grad, gold["dx"]) < 1e-4
    keys = [str(k) for k in gold["grad_keys"]]
    norms = np.array([params[k].grad.norm().item() for k in keys])
    np.testing.assert_allclose(norms, gold["grad_norms"], rtol=2e-4, atol=1e-7)
    for f in gold.files:
        if f.startswith("grad:"):
            assert rel_l2(params[f[5:]].grad, gold[f]) < 1e-4, f
        if f.startswith("after:"):
            np.testing.assert_allclose(work[f[6:]].detach().numpy(), gold[f], rtol=1e-5, atol=1e-6)
    # eval-mode forward with the updated running stats, then the integer Dice path
    with torch.no_grad():
        le = O.unet_forward({k: v.detach() for k, v in work.items()}, x.detach(), bil, training=False)
    assert rel_l2(le, gold["logits_eval"]) < 1e-5
    cnt = O.eval_counts(gold["logits_eval"], y.numpy())
    d = O.dice_from_counts(cnt["inter"], cnt["psum"], cnt["tsum"])
    assert abs(d - float(gold["eval_dice"])) < 1e-6


def test_dice_kats(golden_dir):
    g = np.load(os.path.join(golden_dir, "dice_kats.npz"))
    a = torch.tensor([[1., 1.], [0., 0.]])
    b = torch.tensor([[1., 0.], [1., 0.]])
    assert O.dice_coeff(torch.ones(1, 2, 2), torch.ones(1, 2, 2)).item() == float(g["ones"]) == 1.0
    assert O.dice_coeff(torch.zeros(1, 2, 2), torch.zeros(1, 2, 2)).item() == float(g["zeros"]) == 1.0
    assert O.dice_coeff(torch.ones(1, 2, 2), torch.zeros(1, 2, 2)).item() == pytest.approx(2.5e-7, rel=1e-5)
    assert O.dice_coeff(a[None], b[None]).item() == float(g["ab"])
    assert O.dice_loss(torch.stack([a, b]), torch.stack([b, b])).item() == float(g["loss_ab_bb"]) == 0.25
    p, t = torch.from_numpy(g["rand_p"]), torch.from_numpy(g["rand_t"])
    assert O.multiclass_dice_coeff(p, t, False).item() == pytest.approx(float(g["rand_mc_false"]), abs=1e-7)
    assert O.multiclass_dice_coeff(p, t, True).item() == pytest.approx(float(g["rand_mc_true"]), abs=1e-7)
    assert O.dice_loss(p, t, True).item() == pytest.approx(float(g["rand_loss_mc"]), abs=1e-7)
    assert O.dice_loss(p[:, 0], t[:, 0], False).item() == pytest.approx(float(g["rand_loss_bin"]), abs=1e-7)
    with pytest.raises(AssertionError):     # batch-1 squeeze precondition, util/dice_score.py:8
        O.dice_loss(torch.ones(4, 4), torch.ones(4, 4))
    for nm in ("ev3", "ev1"):
        cnt = O.eval_counts(g[nm + "_logits"], g[nm + "_y"])
        d = O.dice_from_counts(cnt["inter"], cnt["psum"], cnt["tsum"])
        assert abs(d - float(g[nm + "_dice"])) < 1e-6, nm
    # tie rule: first index wins
    cnt = O.eval_counts(g["ev3_logits"], g["ev3_y"])
    assert (cnt["mask"][0, 0, :5] == 0).all()


def _unpack_masks(gold):
    ms = []
    for d in (4, 5, 6):
        shp = tuple(int(s) for s in gold[f"dropmask_depth{d}_shape"])
        n = int(np.prod(shp))
        ms.append(torch.from_numpy(np.unpackbits(gold[f"dropmask_depth{d}"])[:n].reshape(shp)).float())
    return ms


@pytest.mark.parametrize("name,g_train,drop", [("pix2pix_train_nodrop", True, False),
                                               ("pix2pix_train_drop", True, True),
                                               ("pix2pix_eval", False, False),
                                               ("pix2pix_unet128_nl2", True, False),
                                               ("pix2pix_unet128_nl4", True, False)])
def test_pix2pix_oracle_matches_reference(name, g_train, drop, golden_dir):
    gold = np.load(os.path.join(golden_dir, name + ".npz"))
    num_downs = int(gold["num_downs"]) if "num_downs" in gold else 8       # unet_256 | unet_128 (networks.py:148-151)
    n_layers = int(gold["n_layers"]) if "n_layers" in gold else 3           # 'basic' | 'n_layers' (networks.py:197-200)
    hw = int(gold["hw"]) if "hw" in gold else 256
    stG = O.seeded_state(O.generator_state_spec(1, 1, 64, num_downs), 11, conv_std=0.02)
    stD = O.seeded_state(O.discriminator_state_spec(2, 64, n_layers), 12, conv_std=0.02)
    arch = torch.from_numpy(gold["arch"]).clone().requires_grad_(True)
    g = torch.Generator().manual_seed(14)
    mask = torch.randint(0, 2, (2, 1, hw, hw), generator=g).float()
    image = torch.rand(2, 1, hw, hw, generator=g)
    assert mask.sum().item() == float(gold["mask_sum"])
    pG = {k: v.clone().requires_grad_(True) for k, v in stG.items()
          if v.dtype.is_floating_point and "running" not in k}
    pD = {k: v.clone().requires_grad_(True) for k, v in stD.items()
          if v.dtype.is_floating_point and "running" not in k}
    wG, wD = dict(stG), dict(stD)
    wG.update(pG)
    wD.update(pD)
    masks = _unpack_masks(gold) if drop else None
    fake = O.generator_forward(wG, mask, arch, training=g_train, dropout_masks=masks, num_downs=num_downs)
    assert rel_l2(fake.detach().flatten()[::13], gold["fake_sample"]) < 2e-5
    pred_fake_d = O.discriminator_forward(wD, torch.cat([mask, fake.detach()], 1))
    pred_real = O.discriminator_forward(wD, torch.cat([mask, image], 1))
    loss_D = 0.5 * (O.gan_loss(pred_fake_d, False) + O.gan_loss(pred_real, True))
    gD = torch.autograd.grad(loss_D, [pD[str(k)] for k in gold["gradD_keys"]])
    pred_fake = O.discriminator_forward(wD, torch.cat([mask, fake], 1))
    g_gan = O.gan_loss(pred_fake, True)
    g_l1 = F.l1_loss(fake, image) * 100.0
    keysG = [str(k) for k in gold["gradG_keys"]]
    gs = torch.autograd.grad(g_gan + g_l1, [pG[k] for k in keysG] + [arch])
    assert abs(loss_D.item() - float(gold["loss_D"])) < 1e-5
    assert abs(g_gan.item() - float(gold["G_GAN"])) < 1e-5
    assert abs(g_l1.item() - float(gold["G_L1"])) < 2e-4
    assert rel_l2(pred_fake.detach(), gold["pred_fake"]) < 1e-4
    assert rel_l2(pred_real.detach(), gold["pred_real"]) < 1e-4
    np.testing.assert_allclose([t.norm().item() for t in gD], gold["gradD_norms"], rtol=1e-3, atol=1e-8)
    np.testing.assert_allclose([t.norm().item() for t in gs[:-1]], gold["gradG_norms"], rtol=2e-3, atol=1e-8)
    assert rel_l2(gs[-1], gold["grad_arch"]) < 1e-3
    assert rel_l2(gs[keysG.index("model.model.0.weight")], gold["gradG:model.model.0.weight"]) < 1e-3


def test_merged_upconv_equals_mixed():
    """the three concentric stride-2 transposed convs collapse into one 8x8."""
    g = torch.Generator().manual_seed(5)
    st = {f"c._ops._ops.{i}.op.weight": torch.randn(6, 5, k, k, generator=g, dtype=torch.float64)
          for i, k in enumerate((4, 6, 8))}
    st.update({f"c._ops._ops.{i}.op.bias": torch.randn(5, generator=g, dtype=torch.float64) for i in range(3)})
    alpha = torch.randn(3, generator=g, dtype=torch.float64)
    x = torch.randn(2, 6, 7, 5, generator=g, dtype=torch.float64)
    ref = O.mixed_upconv(x, st, "c", alpha, bias=True)
    W, b = O.merged_upconv_weight(st, "c", alpha, bias=True)
    out = F.conv_transpose2d(x, W, b, stride=2, padding=3)
    assert (ref - out).abs().max().item() < 1e-12


def _protocol_a_state():
    """SURVEY 8c Protocol A: the reference's default initialisation under torch.manual_seed(1234), drawn through plain
    torch.nn containers in the reference's construction order (what both UNet classes do)."""
    from semantic_segmentation_b200.unet import UNet           # parameter containers only: no CUDA needed to construct
    torch.manual_seed(1234)
    return {k: v.clone() for k, v in UNet(1, 2, True).state_dict().items()}


def test_oracle_on_protocol_a(golden_dir):
    """BASELINE config 1 (UNet(1,2,bilinear) on 4x1x256x256): the oracle against the reference's own run -- logits mean
    0.155573, CE 0.724931, the bit-exact counts 108265 / 53905 / 239025, every convolution output and a sample of every
    parameter gradient."""
    gold = np.load(os.path.join(golden_dir, "unet_protocol_a.npz"))
    st = _protocol_a_state()
    keys = [str(k) for k in gold["state_keys"]]
    np.testing.assert_allclose([st[k].double().sum().item() for k in keys], gold["state_sums"], rtol=0, atol=1e-9)
    g = torch.Generator().manual_seed(4321)
    x = torch.rand(4, 1, 256, 256, generator=g)
    y = torch.randint(0, 2, (4, 256, 256), generator=g)
    params = {k: v.clone().requires_grad_(True) for k, v in st.items() if v.dtype.is_floating_point and "running" not in k}
    work = dict(st)
    work.update(params)
    taps = {}
    logits = O.unet_forward(work, x, True, training=True, taps=taps)
    loss, ce, dl = O.seg_loss(logits, y)
    loss.backward()
    assert rel_l2(logits.detach(), gold["logits"]) < 1e-5
    assert abs(logits.mean().item() - 0.155573) < 1e-6 and abs(ce.item() - 0.724931) < 1e-6
    assert abs(dl.item() - 0.500099) < 1e-6
    c = O.eval_counts(gold["logits"], y.numpy())
    assert (int(c["psum"].sum()), int(c["inter"].sum()), int(c["psum"].sum() + c["tsum"].sum())) == (108265, 53905, 239025)

    def take(t):
        f = t.detach().reshape(-1)
        s = max(1, f.numel() // 2048)
        return f[::(s | 1 if s > 1 else 1)]

    for k in gold["tap_keys"]:
        k = str(k)
        if k == "outc.conv":
            continue
        assert rel_l2(take(taps[k]), gold["tap:" + k]) < 1e-5, k
    # gradients: within 3x of the reference's own fp32-vs-fp64 distance (ReLU-branch flips), floor 1e-5
    for k in gold["grad_keys"]:
        k = str(k)
        e = rel_l2(take(params[k].grad), gold["grad64:" + k])
        e_ref = rel_l2(gold["grad:" + k], gold["grad64:" + k])
        assert e <= max(3 * e_ref, 1e-5), (k, e, e_ref)
