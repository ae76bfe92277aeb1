"""The Betty-compatible bilevel driver over the CUDA modules: the GenSeg end-to-end loop of
running_files/train_end2end_jsrt.py:141-342 (four problems, the script's dependency graph, roll_back) and the DARTS
finite-difference hyper-gradient through the CUDA generator (in-place ``p.data`` perturbation, ``autograd.grad`` w.r.t.
``networks.upconv_arch``).  betty-ml's numerics are unpinned by the reference; the closed-form checks of the driver itself
are in tests/test_betty_driver.py (CPU)."""
import math

import pytest
import torch
from torch import nn

pytestmark = pytest.mark.gpu

import semantic_segmentation_b200 as S  # noqa: E402
from oracle import genseg_oracle as O  # noqa: E402
from semantic_segmentation_b200.betty.configs import Config, EngineConfig  # noqa: E402
from semantic_segmentation_b200.betty.engine import Engine  # noqa: E402
from semantic_segmentation_b200.betty.problems import ImplicitProblem  # noqa: E402
from semantic_segmentation_b200.models_pix2pix import networks  # noqa: E402
from semantic_segmentation_b200.optim import Adam, RMSprop  # noqa: E402
from semantic_segmentation_b200.unet import UNet  # noqa: E402
from semantic_segmentation_b200.unet.evaluate import evaluate  # noqa: E402
from semantic_segmentation_b200.util import augment  # noqa: E402
from semantic_segmentation_b200.util.dice_score import seg_loss  # noqa: E402

DEV = "cuda"


def _lung_masks(batch):
    yy, xx = torch.meshgrid(torch.arange(256), torch.arange(256), indexing="ij")
    lungs = ((((xx - 85) / 50.0) ** 2 + ((yy - 128) / 80.0) ** 2 < 1) | (((xx - 171) / 45.0) ** 2 + ((yy - 130) / 85.0) ** 2 < 1))
    return lungs.float()[None, None].repeat(batch, 1, 1, 1)


def _nets():
    G = networks.define_G(1, 1, 64, "unet_256", "batch", True, "normal", 0.02, [])
    D = networks.define_D(2, 64, "basic", 3, "batch", "normal", 0.02, [])
    G.load_state_dict(O.seeded_state(O.generator_state_spec(1, 1, 64), 11, conv_std=0.02), strict=True)
    D.load_state_dict(O.seeded_state(O.discriminator_state_spec(2, 64), 12, conv_std=0.02), strict=True)
    net = UNet(1, 1, False)
    net.load_state_dict(O.seeded_state(O.unet_state_spec(1, 1, False), 1234))
    return G.to(DEV).train(), D.to(DEV).train(), net.to(DEV).train()


def test_genseg_end_to_end_loop_reference_graph():
    S.set_precision("bf16")
    old_arch = networks.upconv_arch
    try:
        B = 2
        G, D, net = _nets()
        networks.upconv_arch = (1e-3 * torch.randn(8, 3, generator=torch.Generator().manual_seed(5))).to(DEV).requires_grad_(True)
        crit = networks.GANLoss("vanilla").to(DEV)
        g = torch.Generator().manual_seed(14)
        loader = [{"image": torch.rand(B, 1, 256, 256, generator=g), "mask": _lung_masks(B),
                   "image_pix2pix": torch.rand(B, 1, 256, 256, generator=g), "mask_pix2pix": _lung_masks(B)} for _ in range(2)]
        val_loader = [{"image": torch.rand(B, 1, 256, 256, generator=g), "mask": _lung_masks(B)}]
        mats = augment.draw_affine(B, 256, 256, torch.Generator().manual_seed(100)).to(DEV)
        losses = {"G": [], "D": [], "U": [], "A": [], "val": []}

        class Generator(ImplicitProblem):           # train_end2end_jsrt.py:141-155
            def training_step(self, batch):
                real_mask, real_image = batch["mask_pix2pix"], batch["image_pix2pix"]
                fake_image = self.module(real_mask)
                pred_fake = self.netD(torch.cat((real_mask, fake_image), 1))
                loss = crit(pred_fake, True) + networks.l1_loss(fake_image, real_image) * 100.0
                losses["G"].append(loss.item())
                return loss

        class Discriminator(ImplicitProblem):       # :158-173
            def training_step(self, batch):
                real_mask, real_image = batch["mask_pix2pix"], batch["image_pix2pix"]
                fake_image = self.netG(real_mask)
                pred_fake = self.module(torch.cat((real_mask, fake_image), 1).detach())
                pred_real = self.module(torch.cat((real_mask, real_image), 1))
                loss = (crit(pred_fake, False) + crit(pred_real, True)) * 0.5
                losses["D"].append(loss.item())
                return loss

        class Unet(ImplicitProblem):                # :177-232, imgaug / torchvision chain on the device
            def training_step(self, batch):
                images, true_masks = batch["image"], batch["mask"]
                loss = seg_loss(net(images), true_masks.long())
                fake_mask = augment.augment_masks(true_masks, mats)
                fake_image = augment.synth_image_postprocess(self.netG(fake_mask).detach())
                fake_loss = seg_loss(net(fake_image), fake_mask.long())
                total = loss + 1.0 * fake_loss
                losses["U"].append(total.item())
                return total

        class Arch(ImplicitProblem):                # :235-241
            def training_step(self, batch):
                loss = seg_loss(self.unet(batch["image"]), batch["mask"].long())
                losses["A"].append(loss.item())
                return loss

        class SSEngine(Engine):                     # :244-253
            @torch.no_grad()
            def validation(self):
                score = evaluate(self.unet.module, val_loader, torch.device(DEV), False)
                losses["val"].append(float(score))
                return {"val_score": float(score)}

        inner, outer = Config(type="darts", unroll_steps=1), Config(retain_graph=True)
        dev = torch.device(DEV)
        netG = Generator(name="netG", module=G, optimizer=Adam(list(G.parameters()), lr=2e-5, betas=(0.5, 0.999)),
                         train_data_loader=loader, config=inner, device=dev)
        netD = Discriminator(name="netD", module=D, optimizer=Adam(list(D.parameters()), lr=2e-5, betas=(0.5, 0.999)),
                             train_data_loader=loader, config=inner, device=dev)
        unet = Unet(name="unet", module=net, optimizer=RMSprop(list(net.parameters()), lr=1e-5, weight_decay=1e-8, momentum=0.9),
                    train_data_loader=loader, config=inner, device=dev)
        arch = Arch(name="arch", module=net, optimizer=torch.optim.Adam([networks.upconv_arch], lr=1e-4, betas=(0.5, 0.999),
                                                                        weight_decay=1e-5),
                    train_data_loader=val_loader, config=outer, device=dev)
        l2u = {netG: [unet], netG: [arch]}          # noqa: F601 -- verbatim (:334)
        u2l = {arch: [netG]}
        g0 = G.model.model[0].weight.detach().clone()
        u0 = net.inc.double_conv[0].weight.detach().clone()
        eng = SSEngine(config=EngineConfig(valid_step=2, train_iters=2, roll_back=True), problems=[netG, netD, unet, arch],
                       dependencies={"l2u": l2u, "u2l": u2l})
        eng.run()
        torch.cuda.synchronize()
        assert (netG.count, netD.count, unet.count, arch.count) == (2, 2, 2, 2)
        # roll_back: netG re-does its step after every arch step -> 2 training_steps per iteration
        assert len(losses["G"]) == 4 and len(losses["D"]) == 2 and len(losses["U"]) == 2 and len(losses["A"]) == 2
        assert all(math.isfinite(v) for vs in losses.values() for v in vs), losses
        assert len(losses["val"]) == 1 and 0.0 <= losses["val"][0] <= 1.0
        assert not torch.equal(g0, G.model.model[0].weight.detach()) and not torch.equal(u0, net.inc.double_conv[0].weight.detach())
        assert all(torch.isfinite(p).all() for m in (G, D, net) for p in m.parameters())
        print("genseg loop losses", {k: [round(v, 4) for v in vs] for k, vs in losses.items()})
    finally:
        networks.upconv_arch = old_arch


class _ArchHolder(nn.Module):
    """The architecture variables as the parameters of a module, so that they are the Arch problem's variables."""

    def __init__(self, alpha):
        super().__init__()
        self.upconv_arch = nn.Parameter(alpha)

    def forward(self):
        return self.upconv_arch


@pytest.mark.parametrize("precision", ["fp32", "bf16"])
def test_darts_hypergradient_through_cuda_generator(precision):
    """Upper variables = networks.upconv_arch, lower = the generator: the implicit term needs two more generator
    forward/backward passes at G +- eps v (perturbed in place through p.data, so the weight re-pack must pick it up) with
    gradients taken w.r.t. upconv_arch only."""
    S.set_precision(precision)
    old_arch = networks.upconv_arch
    try:
        G, _, _ = _nets()
        G._program.use_dropout = False
        for L in (4, 5, 6):
            G._program.up[L].dropout = False
        holder = _ArchHolder((1e-3 * torch.randn(8, 3, generator=torch.Generator().manual_seed(5))).to(DEV))
        networks.upconv_arch = holder.upconv_arch
        g = torch.Generator().manual_seed(3)
        train = [{"mask": _lung_masks(1), "image": torch.rand(1, 1, 256, 256, generator=g)}]
        val = [{"mask": _lung_masks(1).flip(3), "image": torch.rand(1, 1, 256, 256, generator=g)}]
        seen = {"implicit": None, "calls": 0}

        class Generator(ImplicitProblem):
            def training_step(self, batch):
                seen["calls"] += 1
                return networks.l1_loss(self.module(batch["mask"]), batch["image"])

        class Arch(ImplicitProblem):
            def training_step(self, batch):
                return networks.l1_loss(self.netG(batch["mask"]), batch["image"])

            def custom_optimizer_step(self):
                seen["total"] = self.module.upconv_arch.grad.detach().clone()

        dev = torch.device(DEV)
        # darts_alpha 1.0: a unit-norm perturbation of the 54 M generator weights (||w|| ~ 150) is far above rounding
        netG = Generator(name="netG", module=G, optimizer=torch.optim.SGD(G.parameters(), lr=0.0), train_data_loader=train,
                         config=Config(type="darts", unroll_steps=1, darts_alpha=1.0), device=dev)
        arch = Arch(name="arch", module=holder, optimizer=None, train_data_loader=val, config=Config(retain_graph=True), device=dev)
        w0 = [p.detach().clone() for p in G.parameters()]
        eng = Engine(config=EngineConfig(train_iters=1, valid_step=0), problems=[netG, arch],
                     dependencies={"l2u": {netG: [arch]}, "u2l": {arch: [netG]}})
        eng.run()
        torch.cuda.synchronize()
        assert seen["calls"] == 3                                  # the step + the two finite-difference evaluations
        total = seen["total"]
        # direct term alone, same batch, same weights (lr 0)
        direct = torch.autograd.grad(networks.l1_loss(G(val[0]["mask"].to(DEV)), val[0]["image"].to(DEV)), holder.upconv_arch)[0]
        implicit = total - direct
        print(precision, "direct", direct.norm().item(), "implicit", implicit.norm().item())
        assert torch.isfinite(total).all() and direct.norm().item() > 0 and implicit.norm().item() > 0
        # weights are back at w (w + eps v - 2 eps v + eps v in fp32)
        worst = max(((a - b.detach()).abs().max() / a.abs().max().clamp_min(1e-12)).item() for a, b in zip(w0, G.parameters()))
        assert worst < 1e-5, worst
    finally:
        networks.upconv_arch = old_arch
        S.set_precision("bf16")
