"""Generate tests/golden/*.npz by running the UNMODIFIED reference modules.

Run in the build container only (needs /root/reference, read-only):

    PYTHONDONTWRITEBYTECODE=1 python tests/golden/make_golden.py            # U-Net, Dice KATs, unet_256 Pix2Pix cases
    PYTHONDONTWRITEBYTECODE=1 python tests/golden/make_golden.py instance   # --norm instance --gan_mode lsgan
    PYTHONDONTWRITEBYTECODE=1 python tests/golden/make_golden.py variants   # --netG unet_128, --netD n_layers (2 / 4)

The reference ships no tests or golden vectors, so parity is pinned on outputs of
the reference itself, on seeded weights/inputs from oracle.genseg_oracle
(``seeded_state`` / ``seeded_inputs``).  The fixtures travel to the GPU box; the
reference does not.
"""
import os
import sys

import numpy as np
import torch
import torch.nn.functional as F

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, "/root/reference")
sys.dont_write_bytecode = True

from oracle import genseg_oracle as O  # noqa: E402

from unet import UNet  # noqa: E402  (reference)
from unet.evaluate import evaluate as ref_evaluate  # noqa: E402
from util.dice_score import dice_coeff, dice_loss, multiclass_dice_coeff  # noqa: E402
from models_pix2pix import networks  # noqa: E402

torch.set_num_threads(8)


def grads_of(model):
    return {k: p.grad.detach().clone() for k, p in model.named_parameters()}


def unet_case(name, n_channels, n_classes, bilinear, n, h, w, wseed, xseed):
    spec = O.unet_state_spec(n_channels, n_classes, bilinear)
    st = O.seeded_state(spec, wseed)
    net = UNet(n_channels, n_classes, bilinear)
    ref_sd = net.state_dict()
    assert list(ref_sd.keys()) == list(spec.keys()), "state spec keys differ from the reference"
    for k in spec:
        assert tuple(ref_sd[k].shape) == tuple(spec[k]), k
    net.load_state_dict(st, strict=True)
    x, y = O.seeded_inputs(xseed, n, n_channels, h, w, n_classes)
    net.train()
    x.requires_grad_(True)
    logits = net(x)
    logits.retain_grad()
    if n_classes == 1:
        ce = torch.nn.BCEWithLogitsLoss()(logits, y.float())
        dl = dice_loss(torch.sigmoid(logits.squeeze(1)), y.float().squeeze(1), multiclass=False)
    else:
        ce = torch.nn.CrossEntropyLoss()(logits, y)
        dl = dice_loss(F.softmax(logits, dim=1).float(), F.one_hot(y, n_classes).permute(0, 3, 1, 2).float(),
                       multiclass=True)
    loss = ce + dl
    loss.backward()
    g = grads_of(net)
    sd_after = net.state_dict()
    out = dict(
        logits=logits.detach().numpy(), ce=ce.item(), dice=dl.item(), loss=loss.item(),
        dlogits=logits.grad.numpy(), dx=x.grad.numpy(),
        grad_norms=np.array([g[k].norm().item() for k in g], np.float64),
        grad_keys=np.array(list(g.keys())),
        x_sum=x.detach().double().sum().item(), y_sum=int(y.sum()),
    )
    for k in ("inc.double_conv.0.weight", "inc.double_conv.1.weight", "inc.double_conv.1.bias",
              "outc.conv.weight", "outc.conv.bias", "up4.conv.double_conv.4.weight",
              "down4.maxpool_conv.1.double_conv.4.bias"):
        out["grad:" + k] = g[k].numpy()
    # a strided sample of one large grad
    out["grad_sample:down4.maxpool_conv.1.double_conv.3.weight"] = \
        g["down4.maxpool_conv.1.double_conv.3.weight"].flatten()[::997].numpy()
    for k in ("inc.double_conv.1.running_mean", "inc.double_conv.1.running_var",
              "down4.maxpool_conv.1.double_conv.4.running_mean", "down4.maxpool_conv.1.double_conv.4.running_var",
              "up4.conv.double_conv.4.running_var", "inc.double_conv.1.num_batches_tracked"):
        out["after:" + k] = sd_after[k].numpy()
    # eval-mode forward (uses the updated running stats) + evaluate() Dice
    net.eval()
    with torch.no_grad():
        le = net(x.detach())
    out["logits_eval"] = le.numpy()
    loader = [dict(image=x.detach(), mask=y)]
    with torch.no_grad():
        ds = ref_evaluate(net, loader, torch.device("cpu"), False)
    out["eval_dice"] = float(ds)
    np.savez_compressed(os.path.join(HERE, name + ".npz"), **out)
    print(name, "loss", loss.item(), "ce", ce.item(), "dice", dl.item(), "eval_dice", float(ds))


def dice_kats():
    out = {}
    a = torch.tensor([[1., 1.], [0., 0.]])
    b = torch.tensor([[1., 0.], [1., 0.]])
    out["ones"] = dice_coeff(torch.ones(1, 2, 2), torch.ones(1, 2, 2)).item()
    out["zeros"] = dice_coeff(torch.zeros(1, 2, 2), torch.zeros(1, 2, 2)).item()
    out["ones_zeros"] = dice_coeff(torch.ones(1, 2, 2), torch.zeros(1, 2, 2)).item()
    out["ab"] = dice_coeff(a[None], b[None]).item()
    out["loss_ab_bb"] = dice_loss(torch.stack([a, b]), torch.stack([b, b])).item()
    g = torch.Generator().manual_seed(77)
    p = torch.rand(3, 4, 16, 12, generator=g)
    t = (torch.rand(3, 4, 16, 12, generator=g) > 0.6).float()
    out["rand_p"] = p.numpy()
    out["rand_t"] = t.numpy()
    out["rand_mc_false"] = multiclass_dice_coeff(p, t, reduce_batch_first=False).item()
    out["rand_mc_true"] = multiclass_dice_coeff(p, t, reduce_batch_first=True).item()
    out["rand_loss_mc"] = dice_loss(p, t, multiclass=True).item()
    out["rand_loss_bin"] = dice_loss(p[:, 0], t[:, 0], multiclass=False).item()
    # evaluate() on given logits through a stub net (n_classes 3 and 1), incl. ties / tiny logits
    logits3 = torch.randn(4, 3, 24, 20, generator=g)
    logits3[0, :, 0, :5] = 0.25          # exact ties -> first index wins
    y3 = torch.randint(0, 3, (4, 24, 20), generator=g)
    y3[3] = 0                             # an image with empty foreground
    logits1 = torch.randn(4, 1, 24, 20, generator=g)
    logits1[0, 0, 0, :6] = torch.tensor([0.0, 1e-8, -1e-8, 6e-8, 1.2e-7, 2.4e-7])
    y1 = torch.randint(0, 2, (4, 1, 24, 20), generator=g)

    class Stub(torch.nn.Module):
        def __init__(self, lg):
            super().__init__()
            self.lg = lg
            self.n_classes = lg.shape[1]

        def forward(self, x):
            return self.lg

    for nm, lg, yy in (("ev3", logits3, y3), ("ev1", logits1, y1)):
        ds = ref_evaluate(Stub(lg), [dict(image=torch.zeros(4, 1, 24, 20), mask=yy)], torch.device("cpu"), False)
        out[nm + "_logits"] = lg.numpy()
        out[nm + "_y"] = yy.numpy()
        out[nm + "_dice"] = float(ds)
    np.savez_compressed(os.path.join(HERE, "dice_kats.npz"), **out)
    print("dice kats", {k: v for k, v in out.items() if not hasattr(v, "shape")})


def pix2pix_case(name, use_dropout, g_train, batch=2, netG="unet_256", n_layers=3, hw=256):
    """netG 'unet_128' (networks.py:148-151, 7 levels) and netD 'n_layers' with n_layers != 3 (networks.py:199-200) are the
    flag-reachable variants of the same modules; their fixtures record netG / n_layers / hw."""
    num_downs = {"unet_256": 8, "unet_128": 7}[netG]
    specG = O.generator_state_spec(1, 1, 64, num_downs)
    specD = O.discriminator_state_spec(2, 64, n_layers)
    stG = O.seeded_state(specG, 11, conv_std=0.02)
    stD = O.seeded_state(specD, 12, conv_std=0.02)
    G = networks.define_G(1, 1, 64, netG, "batch", use_dropout, "normal", 0.02, [])
    D = networks.define_D(2, 64, "basic" if n_layers == 3 else "n_layers", n_layers, "batch", "normal", 0.02, [])
    assert list(G.state_dict().keys()) == list(specG.keys())
    assert list(D.state_dict().keys()) == list(specD.keys())
    for k in specG:
        assert tuple(G.state_dict()[k].shape) == tuple(specG[k]), k
    G.load_state_dict(stG, strict=True)
    D.load_state_dict(stD, strict=True)
    ga = torch.Generator().manual_seed(13)
    arch = 0.5 * torch.randn(8, 3, generator=ga)   # larger than the 1e-3 init so the softmax matters
    networks.upconv_arch.data.copy_(arch)
    if networks.upconv_arch.grad is not None:
        networks.upconv_arch.grad = None
    g = torch.Generator().manual_seed(14)
    mask = torch.randint(0, 2, (batch, 1, hw, hw), generator=g).float()
    image = torch.rand(batch, 1, hw, hw, generator=g)
    G.train(g_train)
    D.train()
    masks = []
    if use_dropout:
        def hook(m, inp, outp):
            masks.append((outp != 0) | (inp[0] == 0))
        for m in G.modules():
            if isinstance(m, torch.nn.Dropout):
                m.register_forward_hook(hook)
        torch.manual_seed(99)
    crit = networks.GANLoss("vanilla")
    fake = G(mask)
    # backward_D (pix2pix_model.py:106-118)
    pred_fake_d = D(torch.cat([mask, fake], 1).detach())
    pred_real = D(torch.cat([mask, image], 1))
    loss_D = 0.5 * (crit(pred_fake_d, False) + crit(pred_real, True))
    loss_D.backward()
    gD = grads_of(D)
    for p in D.parameters():
        p.grad = None
    # backward_G (pix2pix_model.py:120-130) with D frozen
    for p in D.parameters():
        p.requires_grad_(False)
    pred_fake = D(torch.cat([mask, fake], 1))
    g_gan = crit(pred_fake, True)
    g_l1 = torch.nn.L1Loss()(fake, image) * 100.0
    (g_gan + g_l1).backward()
    gG = grads_of(G)
    out = dict(
        arch=arch.numpy(), fake_sample=fake.detach().flatten()[::13].numpy(),
        fake_mean=fake.mean().item(), fake_std=fake.std().item(),
        pred_fake=pred_fake.detach().numpy(), pred_real=pred_real.detach().numpy(),
        loss_D=loss_D.item(), G_GAN=g_gan.item(), G_L1=g_l1.item(),
        gradG_norms=np.array([gG[k].norm().item() for k in gG], np.float64), gradG_keys=np.array(list(gG.keys())),
        gradD_norms=np.array([gD[k].norm().item() for k in gD], np.float64), gradD_keys=np.array(list(gD.keys())),
        grad_arch=networks.upconv_arch.grad.numpy().copy(),
        mask_sum=mask.sum().item(), image_sum=image.double().sum().item(),
        num_downs=np.array(num_downs), n_layers=np.array(n_layers), hw=np.array(hw),
    )
    out["gradG:model.model.0.weight"] = gG["model.model.0.weight"].numpy()
    out["gradG:outer_bias8"] = gG["model.model.3._ops._ops.2.op.bias"].numpy()
    out["gradG:outer_w4"] = gG["model.model.3._ops._ops.0.op.weight"].numpy()
    out["gradD:model.0.weight"] = gD["model.0.weight"].numpy()
    out["gradD:model.11.weight"] = gD["model.%d.weight" % (2 + 3 * n_layers)].numpy()      # the 1-channel head
    if use_dropout and g_train:
        # hooks fire innermost-dropout first: blocks with layer_index 1,2,3 -> depth 6,5,4
        assert len(masks) == 3
        for d, m in zip((6, 5, 4), masks):
            out[f"dropmask_depth{d}"] = np.packbits(m.numpy().reshape(-1))
            out[f"dropmask_depth{d}_shape"] = np.array(m.shape)
    np.savez_compressed(os.path.join(HERE, name + ".npz"), **out)
    print(name, "loss_D", loss_D.item(), "G_GAN", g_gan.item(), "G_L1", g_l1.item(), "fake mean/std",
          out["fake_mean"], out["fake_std"])


def pix2pix_instance_case(name="pix2pix_instance", batch=2):
    """--norm instance (networks.py:34-35): InstanceNorm2d generator + discriminator, one D step and one G step.
    Weights: oracle.seeded_state over the reference modules' own state_dict layout (biases on every convolution)."""
    from collections import OrderedDict
    G = networks.define_G(1, 1, 64, "unet_256", "instance", False, "normal", 0.02, [])
    D = networks.define_D(2, 64, "basic", 3, "instance", "normal", 0.02, [])
    G.load_state_dict(O.seeded_state(OrderedDict((k, tuple(v.shape)) for k, v in G.state_dict().items()), 21, conv_std=0.02))
    D.load_state_dict(O.seeded_state(OrderedDict((k, tuple(v.shape)) for k, v in D.state_dict().items()), 22, conv_std=0.02))
    arch = 0.5 * torch.randn(8, 3, generator=torch.Generator().manual_seed(23))
    networks.upconv_arch.data.copy_(arch)
    networks.upconv_arch.grad = None
    g = torch.Generator().manual_seed(24)
    mask = torch.randint(0, 2, (batch, 1, 256, 256), generator=g).float()
    image = torch.rand(batch, 1, 256, 256, generator=g)
    G.train(); D.train()
    crit = networks.GANLoss("lsgan")           # the other flag-reachable loss (networks.py:237-238)
    fake = G(mask)
    pred_fake_d = D(torch.cat([mask, fake], 1).detach())
    pred_real = D(torch.cat([mask, image], 1))
    loss_D = 0.5 * (crit(pred_fake_d, False) + crit(pred_real, True))
    loss_D.backward()
    gD = grads_of(D)
    for p in D.parameters():
        p.grad = None
        p.requires_grad_(False)
    pred_fake = D(torch.cat([mask, fake], 1))
    g_gan = crit(pred_fake, True)
    g_l1 = torch.nn.L1Loss()(fake, image) * 100.0
    (g_gan + g_l1).backward()
    gG = grads_of(G)
    out = dict(arch=arch.numpy(), fake_sample=fake.detach().flatten()[::13].numpy(), pred_fake=pred_fake.detach().numpy(),
               pred_real=pred_real.detach().numpy(), loss_D=loss_D.item(), G_GAN=g_gan.item(), G_L1=g_l1.item(),
               gradG_norms=np.array([gG[k].norm().item() for k in gG], np.float64), gradG_keys=np.array(list(gG.keys())),
               gradD_norms=np.array([gD[k].norm().item() for k in gD], np.float64), gradD_keys=np.array(list(gD.keys())),
               grad_arch=networks.upconv_arch.grad.numpy().copy())
    out["gradG:model.model.0.weight"] = gG["model.model.0.weight"].numpy()
    out["gradG:outer_w4"] = gG["model.model.3._ops._ops.0.op.weight"].numpy()
    out["gradD:model.0.weight"] = gD["model.0.weight"].numpy()
    out["gradD:model.11.weight"] = gD["model.11.weight"].numpy()      # the 1-channel head of the 'basic' (n_layers 3) net
    np.savez_compressed(os.path.join(HERE, name + ".npz"), **out)
    print(name, "loss_D", loss_D.item(), "G_GAN", g_gan.item(), "G_L1", g_l1.item())


if __name__ == "__main__":
    if len(sys.argv) > 1 and sys.argv[1] == "instance":
        pix2pix_instance_case()
        sys.exit(0)
    if len(sys.argv) > 1 and sys.argv[1] == "variants":
        pix2pix_case("pix2pix_unet128_nl2", use_dropout=False, g_train=True, netG="unet_128", n_layers=2, hw=128)
        pix2pix_case("pix2pix_unet128_nl4", use_dropout=False, g_train=True, netG="unet_128", n_layers=4, hw=256)
        sys.exit(0)
    unet_case("unet_c2_bilinear_32", 1, 2, True, 2, 32, 32, wseed=1, xseed=2)
    unet_case("unet_c1_convT_32", 1, 1, False, 2, 32, 32, wseed=3, xseed=4)
    unet_case("unet_c2_bilinear_odd", 1, 2, True, 2, 36, 44, wseed=5, xseed=6)
    unet_case("unet_c3_rgb_convT_48", 3, 3, False, 3, 48, 32, wseed=7, xseed=8)
    dice_kats()
    pix2pix_case("pix2pix_train_nodrop", use_dropout=False, g_train=True)
    pix2pix_case("pix2pix_train_drop", use_dropout=True, g_train=True)
    pix2pix_case("pix2pix_eval", use_dropout=True, g_train=False)
