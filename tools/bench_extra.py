#!/usr/bin/env python
"""Secondary workloads of BASELINE.json (not the headline bench.py line):

  config 3  Pix2Pix unet_256 generator + 70x70 PatchGAN training step (GANLoss + 100*L1), `optimize_parameters`
            of models_pix2pix/pix2pix_model.py:132-143 restated with our modules: G fwd, D step on (fake.detach, real),
            G step with D frozen; Adam(betas=(0.5, 0.999)) on both.
  config 5  U-Net inference sweep (unet/evaluate.py path): eval forward + fused argmax / Dice counts at
            (N, H) = (64,256), (16,512), (4,1024) and batch-1 latency.

One JSON line per measurement; CUDA-event timing after warm-up.  Usage: python tools/bench_extra.py [--batch 8]
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402


def timed(fn, steps, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.profiler.start()   # ncu --profile-from-start off sees exactly the timed steps
    s.record()
    for _ in range(steps):
        fn()
    e.record()
    torch.cuda.synchronize()
    torch.cuda.profiler.stop()
    return s.elapsed_time(e) * 1e-3 / steps


def pix2pix(batch, steps):
    import semantic_segmentation_b200 as S
    from oracle import genseg_oracle as O
    from semantic_segmentation_b200.models_pix2pix import networks
    from semantic_segmentation_b200.optim import Adam
    S.set_precision("bf16")
    dev = "cuda"
    G = networks.define_G(1, 1, 64, "unet_256", "batch", True, "normal", 0.02, [])
    D = networks.define_D(2, 64, "basic", 3, "batch", "normal", 0.02, [])
    G.load_state_dict(O.seeded_state(O.generator_state_spec(1, 1, 64), 11, conv_std=0.02), strict=True)
    D.load_state_dict(O.seeded_state(O.discriminator_state_spec(2, 64), 12, conv_std=0.02), strict=True)
    G, D = G.to(dev).train(), D.to(dev).train()
    networks.upconv_arch = (1e-3 * torch.randn(8, 3, generator=torch.Generator().manual_seed(5))).to(dev).requires_grad_(True)
    crit = networks.GANLoss("vanilla").to(dev)
    optG = Adam(list(G.parameters()), lr=2e-5, betas=(0.5, 0.999))
    optD = Adam(list(D.parameters()), lr=2e-5, betas=(0.5, 0.999))
    g = torch.Generator().manual_seed(14)
    mask = torch.randint(0, 2, (batch, 1, 256, 256), generator=g).float().to(dev)
    image = torch.rand(batch, 1, 256, 256, generator=g).to(dev)

    def step():
        fake = G(mask)
        for p in D.parameters():
            p.requires_grad_(True)
        optD.zero_grad()
        loss_D = 0.5 * (crit(D(torch.cat([mask, fake], 1).detach()), False) + crit(D(torch.cat([mask, image], 1)), True))
        loss_D.backward()
        optD.step()
        for p in D.parameters():
            p.requires_grad_(False)
        optG.zero_grad()
        loss_G = crit(D(torch.cat([mask, fake], 1)), True) + networks.l1_loss(fake, image) * 100.0
        loss_G.backward()
        optG.step()
        return loss_G

    t = timed(step, steps)
    # algorithmic FLOPs per image: merged-kernel form ~157 GF, as the reference executes it ~234 GF
    print(json.dumps({"workload": "pix2pix optimize_parameters (unet_256 G + basic D, GANLoss + 100*L1, Adam)", "batch": batch,
                      "ms_per_step": t * 1e3, "images_per_s": batch / t, "tflops_merged_form": 157e9 * batch / t / 1e12,
                      "dtype": "bf16", "data": "synthetic"}), flush=True)
    if os.environ.get("P2P_GRAPH", "0") == "1":
        # Diagnostic only: is the eager step (~510 launches through ctypes) host-bound?  Replays one captured step; the
        # Adam step count and the dropout seed are frozen in the graph, so this is a timing probe, not a training loop.
        try:
            side = torch.cuda.Stream()
            side.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(side):
                step()
            torch.cuda.current_stream().wait_stream(side)
            torch.cuda.synchronize()
            graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(graph):
                step()
            tg = timed(graph.replay, steps)
            print(json.dumps({"workload": "pix2pix step, CUDA-graph replay (diagnostic)", "batch": batch, "ms_per_step": tg * 1e3,
                              "images_per_s": batch / tg}), flush=True)
        except Exception as exc:  # noqa: BLE001
            print(json.dumps({"workload": "pix2pix step, CUDA-graph replay (diagnostic)", "error": repr(exc)[:300]}), flush=True)


def evaluate_sweep(steps):
    import semantic_segmentation_b200 as S
    from oracle import genseg_oracle as O
    from semantic_segmentation_b200.unet import UNet
    from semantic_segmentation_b200.unet.evaluate import dice_from_counts, predict_counts
    S.set_precision("bf16")
    dev = "cuda"
    net = UNet(1, 2, True)
    net.load_state_dict(O.seeded_state(O.unet_state_spec(1, 2, True), 1234))
    net = net.to(dev).eval()
    for n, hw in [(64, 256), (16, 512), (4, 1024), (1, 256), (1, 512), (1, 1024)]:
        g = torch.Generator().manual_seed(3)
        x = torch.rand(n, 1, hw, hw, generator=g).to(dev)
        y = torch.randint(0, 2, (n, hw, hw), generator=g).to(dev)

        def step():
            with torch.no_grad():
                counts, _ = predict_counts(net, x, y)
                return dice_from_counts(counts)

        t = timed(step, steps)
        print(json.dumps({"workload": "U-Net(1,2,bilinear) eval forward + fused argmax/Dice counts", "batch": n, "hw": hw,
                          "ms_per_batch": t * 1e3, "images_per_s": n / t, "mpixel_per_s": n * hw * hw / t / 1e6,
                          "fwd_tflops": 79.818e9 * (hw / 256.0) ** 2 * n / t / 1e12, "dtype": "bf16"}), flush=True)


def genseg_iteration(batch, steps):
    """BASELINE config 3/4 -- one iteration of the GenSeg end-to-end loop of running_files/train_end2end_jsrt.py:141-236 with
    our modules: Generator.training_step (:141-155) + Adam, Discriminator.training_step (:158-173) + Adam,
    Unet.training_step (:177-232: real batch, device-side mask augmentation -> G -> image chain -> generated batch,
    unet_loss = loss + loss_lambda * fake_loss) + RMSprop.  UNet(1, 1, bilinear=False), GANLoss('vanilla'), lambda_L1 100,
    loss_lambda 1.0 (scripts/train_end2end_jsrt.sh).  NOT included: the Arch problem's hyper-gradient -- it is computed by
    the external betty-ml engine (DARTS finite differences), absent offline and unpinned by the reference; its
    training_step (one U-Net forward + loss on a validation batch) is timed as forward + backward to the input."""
    import semantic_segmentation_b200 as S
    from oracle import genseg_oracle as O
    from semantic_segmentation_b200.models_pix2pix import networks
    from semantic_segmentation_b200.optim import Adam, RMSprop
    from semantic_segmentation_b200.unet import UNet
    from semantic_segmentation_b200.util import augment
    from semantic_segmentation_b200.util.dice_score import seg_loss
    S.set_precision("bf16")
    dev = "cuda"
    G = networks.define_G(1, 1, 64, "unet_256", "batch", True, "normal", 0.02, [])
    D = networks.define_D(2, 64, "basic", 3, "batch", "normal", 0.02, [])
    G.load_state_dict(O.seeded_state(O.generator_state_spec(1, 1, 64), 11, conv_std=0.02), strict=True)
    D.load_state_dict(O.seeded_state(O.discriminator_state_spec(2, 64), 12, conv_std=0.02), strict=True)
    net = UNet(1, 1, False)
    net.load_state_dict(O.seeded_state(O.unet_state_spec(1, 1, False), 1234))
    G, D, net = G.to(dev).train(), D.to(dev).train(), net.to(dev).train()
    networks.upconv_arch = (1e-3 * torch.randn(8, 3, generator=torch.Generator().manual_seed(5))).to(dev).requires_grad_(True)
    crit = networks.GANLoss("vanilla").to(dev)
    optG = Adam(list(G.parameters()), lr=2e-5, betas=(0.5, 0.999))
    optD = Adam(list(D.parameters()), lr=2e-5, betas=(0.5, 0.999))
    optU = RMSprop(list(net.parameters()), lr=1e-5, weight_decay=1e-8, momentum=0.9)
    g = torch.Generator().manual_seed(14)
    yy, xx = torch.meshgrid(torch.arange(256), torch.arange(256), indexing="ij")
    lungs = ((((xx - 85) / 50.0) ** 2 + ((yy - 128) / 80.0) ** 2 < 1) | (((xx - 171) / 45.0) ** 2 + ((yy - 130) / 85.0) ** 2 < 1))
    mask = lungs.float()[None, None].repeat(batch, 1, 1, 1).to(dev)               # JSRT-shaped lung masks
    image = torch.rand(batch, 1, 256, 256, generator=g).to(dev)
    val_image, val_mask = torch.rand(batch, 1, 256, 256, generator=g).to(dev), mask.clone()
    mats = [augment.draw_affine(batch, 256, 256, torch.Generator().manual_seed(100 + i)).to(dev) for i in range(8)]
    state = {"i": 0}

    def set_d_grad(flag):
        for p in D.parameters():
            p.requires_grad_(flag)

    def step():
        # --- Generator problem
        set_d_grad(False)
        optG.zero_grad()
        fake = G(mask)
        loss_G = crit(D(torch.cat([mask, fake], 1)), True) + networks.l1_loss(fake, image) * 100.0
        loss_G.backward()
        optG.step()
        # --- Discriminator problem
        set_d_grad(True)
        optD.zero_grad()
        with torch.no_grad():
            fake = G(mask)
        loss_D = 0.5 * (crit(D(torch.cat([mask, fake], 1)), False) + crit(D(torch.cat([mask, image], 1)), True))
        loss_D.backward()
        optD.step()
        # --- Unet problem: real batch + generated batch
        optU.zero_grad()
        loss = seg_loss(net(image), mask.long())
        fake_mask = augment.augment_masks(mask, mats[state["i"] % len(mats)])
        state["i"] += 1
        with torch.no_grad():
            fake_image = augment.synth_image_postprocess(G(fake_mask))
        fake_loss = seg_loss(net(fake_image), fake_mask.long())
        (loss + 1.0 * fake_loss).backward()
        optU.step()
        # --- Arch problem's training_step on the validation batch (loss + gradient; hyper-gradient not included)
        vi = val_image.clone().requires_grad_(True)
        seg_loss(net(vi), val_mask.long()).backward()
        return loss_G

    t = timed(step, steps)
    print(json.dumps({"workload": "GenSeg end-to-end iteration (G step + D step + U-Net step on real + generated batch + Arch "
                                  "training_step; Betty hyper-gradient not included)", "batch": batch, "ms_per_iteration": t * 1e3,
                      "real_images_per_s": batch / t, "dtype": "bf16", "data": "synthetic lung-shaped masks"}), flush=True)


def genseg_betty(batch, steps, intended=False):
    """BASELINE config 4 through the Betty-compatible driver (semantic_segmentation_b200/betty): the four
    problems and the dependency graph of train_end2end_jsrt.py:141-342 (``l2u = {netG: [arch]}`` after Python's
    repeated-key rule, ``u2l = {arch: [netG]}``, inner Config(type='darts', unroll_steps=1), roll_back=True); one
    ``Engine.train_step`` = one iteration.  With the script's graph the Arch loss does not reach the generator, so the
    DARTS term vanishes and only its bookkeeping (roll-back snapshot of G + Adam state, G's re-done step) is paid.
    ``intended=True`` times the graph the method describes instead: Arch owns ``upconv_arch`` and its validation loss
    runs through the generator, so every iteration carries the two finite-difference generator passes."""
    import semantic_segmentation_b200 as S
    from oracle import genseg_oracle as O
    from semantic_segmentation_b200.betty import Config, Engine, EngineConfig, ImplicitProblem
    from semantic_segmentation_b200.models_pix2pix import networks
    from semantic_segmentation_b200.optim import Adam, RMSprop
    from semantic_segmentation_b200.unet import UNet
    from semantic_segmentation_b200.util import augment
    from semantic_segmentation_b200.util.dice_score import seg_loss
    S.set_precision("bf16")
    dev = torch.device("cuda")
    G = networks.define_G(1, 1, 64, "unet_256", "batch", True, "normal", 0.02, [])
    D = networks.define_D(2, 64, "basic", 3, "batch", "normal", 0.02, [])
    G.load_state_dict(O.seeded_state(O.generator_state_spec(1, 1, 64), 11, conv_std=0.02), strict=True)
    D.load_state_dict(O.seeded_state(O.discriminator_state_spec(2, 64), 12, conv_std=0.02), strict=True)
    net = UNet(1, 1, False)
    net.load_state_dict(O.seeded_state(O.unet_state_spec(1, 1, False), 1234))
    G, D, net = G.to(dev).train(), D.to(dev).train(), net.to(dev).train()

    class ArchVars(torch.nn.Module):
        def __init__(self):
            super().__init__()
            self.upconv_arch = torch.nn.Parameter((1e-3 * torch.randn(8, 3, generator=torch.Generator().manual_seed(5))).to(dev))

    arch_vars = ArchVars()
    networks.upconv_arch = arch_vars.upconv_arch
    crit = networks.GANLoss("vanilla").to(dev)
    g = torch.Generator().manual_seed(14)
    yy, xx = torch.meshgrid(torch.arange(256), torch.arange(256), indexing="ij")
    lungs = ((((xx - 85) / 50.0) ** 2 + ((yy - 128) / 80.0) ** 2 < 1) | (((xx - 171) / 45.0) ** 2 + ((yy - 130) / 85.0) ** 2 < 1))
    mask = lungs.float()[None, None].repeat(batch, 1, 1, 1)
    # batches live in pinned host memory; the driver moves them to the device every step (Problem.get_batch)
    loader = [{"image": torch.rand(batch, 1, 256, 256, generator=g).pin_memory(), "mask": mask.clone().pin_memory(),
               "image_pix2pix": torch.rand(batch, 1, 256, 256, generator=g).pin_memory(), "mask_pix2pix": mask.clone().pin_memory()}
              for _ in range(4)]
    val_loader = [{"image": torch.rand(batch, 1, 256, 256, generator=g).pin_memory(), "mask": mask.clone().pin_memory()}]
    mats = [augment.draw_affine(batch, 256, 256, torch.Generator().manual_seed(100 + i)).to(dev) for i in range(8)]
    state = {"i": 0}

    class Generator(ImplicitProblem):
        def training_step(self, batch):
            fake = self.module(batch["mask_pix2pix"])
            return crit(self.netD(torch.cat((batch["mask_pix2pix"], fake), 1)), True) + networks.l1_loss(fake, batch["image_pix2pix"]) * 100.0

    class Discriminator(ImplicitProblem):
        def training_step(self, batch):
            with torch.no_grad():
                fake = self.netG(batch["mask_pix2pix"])
            return 0.5 * (crit(self.module(torch.cat((batch["mask_pix2pix"], fake), 1)), False)
                          + crit(self.module(torch.cat((batch["mask_pix2pix"], batch["image_pix2pix"]), 1)), True))

    class Unet(ImplicitProblem):
        def training_step(self, batch):
            loss = seg_loss(self.module(batch["image"]), batch["mask"].long())
            fake_mask = augment.augment_masks(batch["mask"], mats[state["i"] % len(mats)])
            state["i"] += 1
            with torch.no_grad():
                fake_image = augment.synth_image_postprocess(self.netG(fake_mask))
            return loss + 1.0 * seg_loss(self.module(fake_image), fake_mask.long())

    class Arch(ImplicitProblem):
        def training_step(self, batch):
            if intended:     # validation loss of the U-Net on images the generator makes from the validation masks
                return seg_loss(self.unet(self.netG(batch["mask"])), batch["mask"].long())
            return seg_loss(self.unet(batch["image"]), batch["mask"].long())

    inner, outer = Config(type="darts", unroll_steps=1), Config(retain_graph=True)
    netG = Generator(name="netG", module=G, optimizer=Adam(list(G.parameters()), lr=2e-5, betas=(0.5, 0.999)),
                     train_data_loader=loader, config=inner, device=dev)
    netD = Discriminator(name="netD", module=D, optimizer=Adam(list(D.parameters()), lr=2e-5, betas=(0.5, 0.999)),
                         train_data_loader=loader, config=inner, device=dev)
    unet = Unet(name="unet", module=net, optimizer=RMSprop(list(net.parameters()), lr=1e-5, weight_decay=1e-8, momentum=0.9),
                train_data_loader=loader, config=inner, device=dev)
    arch = Arch(name="arch", module=arch_vars if intended else net,
                optimizer=Adam([arch_vars.upconv_arch], lr=1e-4, betas=(0.5, 0.999), weight_decay=1e-5),
                train_data_loader=val_loader, config=outer, device=dev)
    eng = Engine(config=EngineConfig(valid_step=0, train_iters=1, roll_back=True), problems=[netG, netD, unet, arch],
                 dependencies={"l2u": {netG: [arch]}, "u2l": {arch: [netG]}})

    def step():
        eng.global_step += 1
        eng.train_step()

    t = timed(step, steps)
    print(json.dumps({"workload": "GenSeg end-to-end iteration through the Betty-compatible driver (%s graph; G, D, U-Net, Arch problems, "
                                  "roll_back, host batches copied every step)" % ("intended: DARTS hyper-gradient through G" if intended else "reference script"),
                      "batch": batch, "ms_per_iteration": t * 1e3, "real_images_per_s": batch / t, "dtype": "bf16",
                      "data": "synthetic lung-shaped masks"}), flush=True)


def augment_chain(steps):
    """the image chain of the end-to-end step (train_end2end_jsrt.py:186-200) at its size (8 x 1 x 256 x 256)
    and at a bandwidth-relevant size (64 x 1 x 1024 x 1024), device-resident, against the CPU restatement."""
    import time
    from oracle import augment_oracle as AO
    from semantic_segmentation_b200.util import augment as A
    for n, hw in [(8, 256), (64, 1024)]:
        g = torch.Generator().manual_seed(1)
        x = torch.tanh(torch.randn(n, 1, hw, hw, generator=g))
        xd = x.cuda()
        t = timed(lambda: A.synth_image_postprocess(xd), steps)
        # algorithmic bytes: fp32 read twice (min/max, quantise) + u8 write + u8 read + fp32 write = 14 B / pixel
        cpu = None
        if n == 8:
            t0 = time.perf_counter(); AO.synth_image_postprocess(x); cpu = time.perf_counter() - t0
        print(json.dumps({"workload": "synth image chain (batch min-max, uint8, equalize, gamma 0.5)", "batch": n, "hw": hw,
                          "us_per_batch": t * 1e6, "mpixel_per_s": n * hw * hw / t / 1e6, "gbytes_per_s": 14.0 * n * hw * hw / t / 1e9,
                          "cpu_oracle_us": cpu * 1e6 if cpu else None}), flush=True)
    m = (torch.rand(8, 1, 256, 256) > 0.7).float().cuda()
    mats = A.draw_affine(8, 256, 256, torch.Generator().manual_seed(2)).cuda()
    t = timed(lambda: A.augment_masks(m, mats), steps)
    print(json.dumps({"workload": "mask affine warp + threshold", "batch": 8, "hw": 256, "us_per_batch": t * 1e6}), flush=True)


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=8)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--only", default="")
    a = ap.parse_args()
    if a.only in ("", "pix2pix"):
        pix2pix(a.batch, a.steps)
    if a.only in ("", "eval"):
        evaluate_sweep(a.steps)
    if a.only in ("", "augment"):
        augment_chain(a.steps)
    if a.only in ("", "e2e"):
        genseg_iteration(a.batch, a.steps)
    if a.only in ("", "betty"):
        genseg_betty(a.batch, a.steps, intended=False)
        genseg_betty(a.batch, a.steps, intended=True)
