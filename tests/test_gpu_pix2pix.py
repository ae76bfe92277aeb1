"""GPU parity of the Pix2Pix path (define_G unet_256 with mixed up-conv cells, define_D basic PatchGAN,
GANLoss, L1) against the fixtures produced by the reference itself (tests/golden/make_golden.py)."""
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

import semantic_segmentation_b200 as S  # noqa: E402
from oracle import genseg_oracle as O  # noqa: E402
from semantic_segmentation_b200.models_pix2pix import networks  # noqa: E402

DEV = "cuda"


def rel(a, b):
    a = torch.as_tensor(a).detach().double().cpu().flatten()
    b = torch.as_tensor(b).detach().double().cpu().flatten()
    return float((a - b).norm() / b.norm().clamp_min(1e-30))


def _masks(gold):
    ms = []
    for d in (4, 5, 6):
        shp = tuple(int(s) for s in gold[f"dropmask_depth{d}_shape"])
        n = int(np.prod(shp))
        m = torch.from_numpy(np.unpackbits(gold[f"dropmask_depth{d}"])[:n].reshape(shp))
        ms.append(m.permute(0, 2, 3, 1).contiguous().to(DEV, torch.uint8))     # NCHW bool -> dense NHWC uint8
    return ms


def run_pix2pix(name, g_train, drop, precision, golden_dir):
    gold = np.load(os.path.join(golden_dir, name + ".npz"))
    S.set_precision(precision)
    old_arch = networks.upconv_arch
    num_downs = int(gold["num_downs"]) if "num_downs" in gold else 8        # unet_256 | unet_128
    n_layers = int(gold["n_layers"]) if "n_layers" in gold else 3            # 'basic' | 'n_layers'
    hw = int(gold["hw"]) if "hw" in gold else 256
    try:
        G = networks.define_G(1, 1, 64, {8: "unet_256", 7: "unet_128"}[num_downs], "batch", num_downs == 8, "normal", 0.02, [])
        D = networks.define_D(2, 64, "basic" if n_layers == 3 else "n_layers", n_layers, "batch", "normal", 0.02, [])
        G.load_state_dict(O.seeded_state(O.generator_state_spec(1, 1, 64, num_downs), 11, conv_std=0.02), strict=True)
        D.load_state_dict(O.seeded_state(O.discriminator_state_spec(2, 64, n_layers), 12, conv_std=0.02), strict=True)
        G, D = G.to(DEV), D.to(DEV)
        networks.upconv_arch = torch.from_numpy(gold["arch"]).to(DEV).requires_grad_(True)   # like pix2pix_model.py:63
        g = torch.Generator().manual_seed(14)
        mask = torch.randint(0, 2, (2, 1, hw, hw), generator=g).float().to(DEV)
        image = torch.rand(2, 1, hw, hw, generator=g).to(DEV)
        G.train(g_train)
        D.train()
        G._dropout_masks = _masks(gold) if drop else None
        if g_train and not drop and num_downs == 8:
            G._program.use_dropout = False
            for L in (4, 5, 6):
                G._program.up[L].dropout = False
        crit = networks.GANLoss("vanilla").to(DEV)
        fake = G(mask)
        pred_fake_d = D(torch.cat([mask, fake], 1).detach())
        pred_real = D(torch.cat([mask, image], 1))
        loss_D = 0.5 * (crit(pred_fake_d, False) + crit(pred_real, True))
        loss_D.backward()
        gD = {k: p.grad.clone() for k, p in D.named_parameters()}
        for p in D.parameters():
            p.grad = None
            p.requires_grad_(False)
        pred_fake = D(torch.cat([mask, fake], 1))
        g_gan = crit(pred_fake, True)
        g_l1 = networks.l1_loss(fake, image) * 100.0
        (g_gan + g_l1).backward()
        torch.cuda.synchronize()
        assert all(p.grad is None for p in D.parameters())        # frozen D: wgrad skipped, dgrad still flows
        gG = {k: p.grad for k, p in G.named_parameters()}
        out = dict(fake=rel(fake.flatten()[::13], gold["fake_sample"]), pred_fake=rel(pred_fake, gold["pred_fake"]),
                   pred_real=rel(pred_real, gold["pred_real"]), loss_D=abs(loss_D.item() - float(gold["loss_D"])),
                   G_GAN=abs(g_gan.item() - float(gold["G_GAN"])), G_L1=abs(g_l1.item() - float(gold["G_L1"])))
        nD = np.array([gD[str(k)].norm().item() for k in gold["gradD_keys"]])
        nG = np.array([gG[str(k)].norm().item() for k in gold["gradG_keys"]])
        out["gradD_norms"] = float(np.max(np.abs(nD - gold["gradD_norms"]) / (gold["gradD_norms"] + 1e-12)))
        out["gradG_norms"] = float(np.max(np.abs(nG - gold["gradG_norms"]) / (gold["gradG_norms"] + 1e-12)))
        out["gradG_norms_median"] = float(np.median(np.abs(nG - gold["gradG_norms"]) / (gold["gradG_norms"] + 1e-12)))
        out["grad_arch"] = rel(networks.upconv_arch.grad, gold["grad_arch"])
        out["gG0"] = rel(gG["model.model.0.weight"], gold["gradG:model.model.0.weight"])
        out["gG_outer_w4"] = rel(gG["model.model.3._ops._ops.0.op.weight"], gold["gradG:outer_w4"])
        out["gG_outer_b8"] = rel(gG["model.model.3._ops._ops.2.op.bias"], gold["gradG:outer_bias8"])
        out["gD0"] = rel(gD["model.0.weight"], gold["gradD:model.0.weight"])
        out["gD11"] = rel(gD["model.%d.weight" % (2 + 3 * n_layers)], gold["gradD:model.11.weight"])   # the 1-channel head
        return out
    finally:
        networks.upconv_arch = old_arch
        S.set_precision("bf16")


CASES = [("pix2pix_train_nodrop", True, False), ("pix2pix_train_drop", True, True), ("pix2pix_eval", False, False)]


@pytest.mark.parametrize("name,g_train,drop", CASES)
def test_pix2pix_fp32_matches_reference(name, g_train, drop, golden_dir):
    """fp32 validation mode.  Forward quantities <= 1e-4 (the 1x1 / 2x2 bottleneck BatchNorms of unet_256 at batch 2
    normalise over 2..8 values, which amplifies fp32 rounding); losses 1e-4; gradients 5e-2 (ReLU-branch flips)."""
    out = run_pix2pix(name, g_train, drop, "fp32", golden_dir)
    print(name, "fp32", out)
    assert out["fake"] < 2e-4 and out["pred_fake"] < 2e-4 and out["pred_real"] < 1e-4, out
    assert out["loss_D"] < 1e-4 and out["G_GAN"] < 1e-4 and out["G_L1"] < 2e-3, out
    assert out["gradD_norms"] < 5e-2 and out["gD0"] < 5e-2 and out["gD11"] < 5e-2, out
    assert out["gradG_norms_median"] < 1e-2 and out["gG_outer_w4"] < 5e-2 and out["gG_outer_b8"] < 5e-2, out
    assert out["grad_arch"] < 5e-2, out


VARIANTS = ["pix2pix_unet128_nl2", "pix2pix_unet128_nl4"]


@pytest.mark.parametrize("name", VARIANTS)
def test_pix2pix_variants_fp32_match_reference(name, golden_dir):
    """Flag-reachable variants of the same modules: --netG unet_128 (7 levels, networks.py:148-151) and
    --netD n_layers with n_layers_D 2 / 4 (networks.py:199-200), at 128^2 and 256^2; same bounds as the unet_256 cases."""
    out = run_pix2pix(name, True, False, "fp32", golden_dir)
    print(name, "fp32", out)
    assert out["fake"] < 2e-4 and out["pred_fake"] < 2e-4 and out["pred_real"] < 1e-4, out
    assert out["loss_D"] < 1e-4 and out["G_GAN"] < 1e-4 and out["G_L1"] < 2e-3, out
    assert out["gradD_norms"] < 5e-2 and out["gD0"] < 5e-2 and out["gD11"] < 5e-2, out
    assert out["gradG_norms_median"] < 1e-2 and out["gG_outer_w4"] < 5e-2 and out["gG_outer_b8"] < 5e-2, out
    assert out["grad_arch"] < 5e-2, out


@pytest.mark.parametrize("name", VARIANTS)
def test_pix2pix_variants_bf16_run_close(name, golden_dir):
    out = run_pix2pix(name, True, False, "bf16", golden_dir)
    print(name, "bf16", out)
    assert out["fake"] < 1e-1 and out["pred_real"] < 5e-2, out
    assert out["loss_D"] < 2e-2 and out["G_GAN"] < 2e-2 and out["G_L1"] < 0.5, out


@pytest.mark.parametrize("name,g_train,drop", CASES[:2])
def test_pix2pix_bf16_runs_close(name, g_train, drop, golden_dir):
    """bf16 production mode (tcgen05 4x4-s2 convs and merged 8x8 transposed convs): losses within 2e-2 of the
    fp32 reference, generator output within 1e-1 rel-L2."""
    out = run_pix2pix(name, g_train, drop, "bf16", golden_dir)
    print(name, "bf16", out)
    assert out["fake"] < 1e-1 and out["pred_real"] < 5e-2, out
    assert out["loss_D"] < 2e-2 and out["G_GAN"] < 2e-2 and out["G_L1"] < 0.5, out


def test_gan_and_l1_losses_match_oracle():
    g = torch.Generator().manual_seed(21)
    p = torch.randn(2, 1, 30, 30, generator=g)
    a, b = torch.rand(2, 1, 64, 64, generator=g), torch.rand(2, 1, 64, 64, generator=g)
    for mode in ("vanilla", "lsgan", "wgangp"):
        for real in (True, False):
            pc = p.clone().requires_grad_(True)
            ref = O.gan_loss(pc, real, mode)
            ref.backward()
            pd = p.to(DEV).requires_grad_(True)
            out = networks.GANLoss(mode).to(DEV)(pd, real)
            (out * 2.0).backward()
            assert abs(out.item() - ref.item()) < 1e-6, (mode, real)
            assert rel(pd.grad, 2.0 * pc.grad) < 1e-5, (mode, real)
    ac = a.clone().requires_grad_(True)
    ref = torch.nn.functional.l1_loss(ac, b)
    ref.backward()
    ad = a.to(DEV).requires_grad_(True)
    out = networks.l1_loss(ad, b.to(DEV))
    out.backward()
    assert abs(out.item() - ref.item()) < 1e-6 and rel(ad.grad, ac.grad) < 1e-6


def _run_instance(precision, golden_dir):
    from collections import OrderedDict
    gold = np.load(os.path.join(golden_dir, "pix2pix_instance.npz"))
    S.set_precision(precision)
    old_arch = networks.upconv_arch
    try:
        G = networks.define_G(1, 1, 64, "unet_256", "instance", False, "normal", 0.02, [])
        D = networks.define_D(2, 64, "basic", 3, "instance", "normal", 0.02, [])
        G.load_state_dict(O.seeded_state(OrderedDict((k, tuple(v.shape)) for k, v in G.state_dict().items()), 21, conv_std=0.02))
        D.load_state_dict(O.seeded_state(OrderedDict((k, tuple(v.shape)) for k, v in D.state_dict().items()), 22, conv_std=0.02))
        assert not any("running" in k for k in G.state_dict()) and "model.2.bias" in D.state_dict()
        G, D = G.to(DEV).train(), D.to(DEV).train()
        networks.upconv_arch = torch.from_numpy(gold["arch"]).to(DEV).requires_grad_(True)
        g = torch.Generator().manual_seed(24)
        mask = torch.randint(0, 2, (2, 1, 256, 256), generator=g).float().to(DEV)
        image = torch.rand(2, 1, 256, 256, generator=g).to(DEV)
        crit = networks.GANLoss("lsgan").to(DEV)
        fake = G(mask)
        pred_fake_d = D(torch.cat([mask, fake], 1).detach())
        pred_real = D(torch.cat([mask, image], 1))
        loss_D = 0.5 * (crit(pred_fake_d, False) + crit(pred_real, True))
        loss_D.backward()
        gD = {k: p.grad.clone() for k, p in D.named_parameters()}
        for p in D.parameters():
            p.grad = None
            p.requires_grad_(False)
        pred_fake = D(torch.cat([mask, fake], 1))
        g_gan = crit(pred_fake, True)
        g_l1 = networks.l1_loss(fake, image) * 100.0
        (g_gan + g_l1).backward()
        torch.cuda.synchronize()
        gG = {k: p.grad for k, p in G.named_parameters()}
        out = dict(fake=rel(fake.flatten()[::13], gold["fake_sample"]), pred_fake=rel(pred_fake, gold["pred_fake"]),
                   pred_real=rel(pred_real, gold["pred_real"]), loss_D=abs(loss_D.item() - float(gold["loss_D"])),
                   G_GAN=abs(g_gan.item() - float(gold["G_GAN"])), G_L1=abs(g_l1.item() - float(gold["G_L1"])))
        # a convolution bias in front of InstanceNorm is cancelled by the mean subtraction: its gradient is pure rounding
        # noise in the reference too -> compare the norms of weight gradients only
        keysG = [str(k) for k in gold["gradG_keys"]]
        sel = [i for i, k in enumerate(keysG) if k.endswith(".weight")]
        nG = np.array([gG[keysG[i]].norm().item() for i in sel])
        out["gradG_norms_median"] = float(np.median(np.abs(nG - gold["gradG_norms"][sel]) / (gold["gradG_norms"][sel] + 1e-12)))
        out["grad_arch"] = rel(networks.upconv_arch.grad, gold["grad_arch"])
        out["gG0"] = rel(gG["model.model.0.weight"], gold["gradG:model.model.0.weight"])
        out["gG_outer_w4"] = rel(gG["model.model.3._ops._ops.0.op.weight"], gold["gradG:outer_w4"])
        out["gD0"] = rel(gD["model.0.weight"], gold["gradD:model.0.weight"])
        out["gD11"] = rel(gD["model.11.weight"], gold["gradD:model.11.weight"])
        return out
    finally:
        networks.upconv_arch = old_arch
        S.set_precision("bf16")


def test_pix2pix_instance_norm_lsgan_fp32_matches_reference(golden_dir):
    """--norm instance --gan_mode lsgan: InstanceNorm2d generator / discriminator with biased convolutions,
    MSE GAN loss; fixture generated by the reference modules (tests/golden/make_golden.py instance)."""
    out = _run_instance("fp32", golden_dir)
    print("instance fp32", out)
    assert out["fake"] < 2e-4 and out["pred_fake"] < 2e-4 and out["pred_real"] < 1e-4, out
    assert out["loss_D"] < 1e-4 and out["G_GAN"] < 1e-4 and out["G_L1"] < 2e-3, out
    assert out["gD0"] < 5e-2 and out["gD11"] < 5e-2 and out["gG_outer_w4"] < 5e-2 and out["gG0"] < 5e-2, out
    assert out["gradG_norms_median"] < 1e-2 and out["grad_arch"] < 5e-2, out


def test_pix2pix_instance_norm_bf16_runs_close(golden_dir):
    out = _run_instance("bf16", golden_dir)
    print("instance bf16", out)
    assert out["fake"] < 1e-1 and out["pred_real"] < 5e-2, out
    assert out["loss_D"] < 2e-2 and out["G_GAN"] < 2e-2 and out["G_L1"] < 0.5, out


def test_pix2pix_step_is_bit_reproducible(golden_dir):
    """Two runs of the bf16 D step + G step (merged up-cells, thin 128 -> 1 output layer, 512 -> 1 head, split-K deep
    levels, dalpha dot products, GAN / L1 reductions): every loss and every gradient bit-identical (no atomics)."""
    a = run_pix2pix_raw("pix2pix_train_nodrop", golden_dir)
    b = run_pix2pix_raw("pix2pix_train_nodrop", golden_dir)
    assert a["losses"] == b["losses"]
    for k in a["gG"]:
        assert torch.equal(a["gG"][k], b["gG"][k]), k
    for k in a["gD"]:
        assert torch.equal(a["gD"][k], b["gD"][k]), k
    assert torch.equal(a["ga"], b["ga"])


def run_pix2pix_raw(name, golden_dir):
    gold = np.load(os.path.join(golden_dir, name + ".npz"))
    S.set_precision("bf16")
    old_arch = networks.upconv_arch
    try:
        G = networks.define_G(1, 1, 64, "unet_256", "batch", False, "normal", 0.02, [])
        D = networks.define_D(2, 64, "basic", 3, "batch", "normal", 0.02, [])
        G.load_state_dict(O.seeded_state(O.generator_state_spec(1, 1, 64, 8), 11, conv_std=0.02), strict=True)
        D.load_state_dict(O.seeded_state(O.discriminator_state_spec(2, 64, 3), 12, conv_std=0.02), strict=True)
        G, D = G.to(DEV).train(), D.to(DEV).train()
        networks.upconv_arch = torch.from_numpy(gold["arch"]).to(DEV).requires_grad_(True)
        g = torch.Generator().manual_seed(14)
        mask = torch.randint(0, 2, (2, 1, 256, 256), generator=g).float().to(DEV)
        image = torch.rand(2, 1, 256, 256, generator=g).to(DEV)
        crit = networks.GANLoss("vanilla").to(DEV)
        fake = G(mask)
        loss_D = 0.5 * (crit(D(torch.cat([mask, fake], 1).detach()), False) + crit(D(torch.cat([mask, image], 1)), True))
        loss_D.backward()
        gD = {k: p.grad.clone() for k, p in D.named_parameters()}
        for p in D.parameters():
            p.grad = None
            p.requires_grad_(False)
        g_gan = crit(D(torch.cat([mask, fake], 1)), True)
        g_l1 = networks.l1_loss(fake, image) * 100.0
        (g_gan + g_l1).backward()
        torch.cuda.synchronize()
        return dict(losses=(loss_D.item(), g_gan.item(), g_l1.item()), gD=gD, gG={k: p.grad.clone() for k, p in G.named_parameters()},
                    ga=networks.upconv_arch.grad.clone())
    finally:
        networks.upconv_arch = old_arch
        S.set_precision("bf16")


def _p2p_trainer(capturable):
    """Pix2Pix optimize_parameters (pix2pix_model.py:132-143) on a seeded unet_256 generator WITH dropout + PatchGAN, Adam
    (lr 2e-4, beta1 0.5) -- `capturable`: Adam's step count and the generator's dropout call count in device memory."""
    from semantic_segmentation_b200 import gan_program, optim
    gan_program.set_capturable(capturable)
    torch.manual_seed(4242)
    G = networks.define_G(1, 1, 64, "unet_256", "batch", True, "normal", 0.02, [])
    D = networks.define_D(2, 64, "basic", 3, "batch", "normal", 0.02, [])
    G.load_state_dict(O.seeded_state(O.generator_state_spec(1, 1, 64, 8), 11, conv_std=0.02), strict=True)
    D.load_state_dict(O.seeded_state(O.discriminator_state_spec(2, 64, 3), 12, conv_std=0.02), strict=True)
    G, D = G.to(DEV).train(), D.to(DEV).train()
    crit = networks.GANLoss("vanilla").to(DEV)
    optG = optim.Adam(list(G.parameters()), lr=2e-4, betas=(0.5, 0.999), capturable=capturable)
    optD = optim.Adam(list(D.parameters()), lr=2e-4, betas=(0.5, 0.999), capturable=capturable)
    g = torch.Generator().manual_seed(14)
    mask = torch.randint(0, 2, (1, 1, 256, 256), generator=g).float().to(DEV)
    image = torch.rand(1, 1, 256, 256, generator=g).to(DEV)
    pD = list(D.parameters())

    def step():
        fake = G(mask)
        for p in pD:
            p.requires_grad_(True)
        optD.zero_grad()
        loss_D = 0.5 * (crit(D(torch.cat([mask, fake], 1).detach()), False) + crit(D(torch.cat([mask, image], 1)), True))
        loss_D.backward()
        optD.step()
        for p in pD:
            p.requires_grad_(False)
        optG.zero_grad()
        loss_G = crit(D(torch.cat([mask, fake], 1)), True) + networks.l1_loss(fake, image) * 100.0
        loss_G.backward()
        optG.step()
        return loss_G.detach()

    return G, D, step


def test_pix2pix_step_replays_as_cuda_graph():
    """The whole optimize_parameters step captured into ONE CUDA graph (capturable Adam: step count in device memory;
    dropout: call count in device memory, explicit masks with the bits of the eager seed rule) and replayed three times
    == five eager steps of the non-capturable path: same dropout masks call for call, same bias corrections."""
    S.set_precision("bf16")
    old_arch = networks.upconv_arch
    from semantic_segmentation_b200 import gan_program
    try:
        networks.upconv_arch = (1e-3 * torch.randn(8, 3, generator=torch.Generator().manual_seed(5))).to(DEV).requires_grad_(True)
        Ga, Da, step_a = _p2p_trainer(False)
        la = [float(step_a()) for _ in range(5)]
        ref = {k: v.detach().clone() for k, v in list(Ga.state_dict().items()) + [("D." + k, v) for k, v in Da.state_dict().items()]}
        networks.upconv_arch = (1e-3 * torch.randn(8, 3, generator=torch.Generator().manual_seed(5))).to(DEV).requires_grad_(True)
        Gb, Db, step_b = _p2p_trainer(True)
        lb = [float(step_b()) for _ in range(2)]                 # eager steps 1, 2 (capturable path)
        torch.cuda.synchronize()
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph):
            out = step_b()                                       # captured, not executed
        for _ in range(3):                                       # steps 3, 4, 5
            graph.replay()
            lb.append(float(out))
        torch.cuda.synchronize()
        got = {k: v.detach().clone() for k, v in list(Gb.state_dict().items()) + [("D." + k, v) for k, v in Db.state_dict().items()]}
        assert la[0] == lb[0], (la, lb)                          # the capturable path draws the very same dropout masks
        for a, b in zip(la, lb):
            assert abs(a - b) <= 1e-4 * max(1.0, abs(a)), (la, lb)
        for k, v in ref.items():
            if v.dtype.is_floating_point:
                err = float((got[k].double() - v.double()).norm() / v.double().norm().clamp_min(1e-30))
                assert err < 1e-5, (k, err)                      # host pow vs device pow in the bias correction: <= 1 ulp of lr
            else:
                assert torch.equal(got[k], v), k                 # num_batches_tracked: 5 forward passes on both sides
    finally:
        gan_program.set_capturable(False)
        networks.upconv_arch = old_arch
        S.set_precision("bf16")
