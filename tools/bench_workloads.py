"""Secondary workloads of BASELINE.json for `bench.py --workload {pix2pix,genseg,eval}` (the default `unet` workload, the
headline, lives in bench.py).  Each builder returns a dict:

    step(batch)        one step of the workload on device-resident inputs, returns a device scalar (a loss)
    host_batch         pinned host tensors of one step's inputs (the end-to-end leg copies them every step)
    to_device(hb)      host batch -> the static device inputs the step reads
    units              images (or real images) processed per step on this rank
    metric, unit, workload, flops_per_unit (algorithmic) ...

config 3  Pix2Pix `optimize_parameters` (models_pix2pix/pix2pix_model.py:132-143): G forward, D step on (fake.detach, real),
          G step with D frozen; unet_256 generator (mixed up-cells, dropout) + 70x70 PatchGAN, GANLoss('vanilla') + 100 * L1,
          Adam(lr 2e-5, betas (0.5, 0.999)); 8 images per GPU (weak scaling: global batch 8 x N).
config 4  one iteration of the GenSeg end-to-end loop (running_files/train_end2end_jsrt.py:141-236): Generator step,
          Discriminator step, U-Net step on the real batch and on a batch generated from device-side augmented masks, and the
          Arch problem's training_step (the DARTS hyper-gradient itself belongs to the external betty-ml engine: the
          Betty-compatible driver is timed by tests / tools, not here); 8 real images per GPU.
config 5  U-Net inference (unet/evaluate.py path): eval-mode forward + fused argmax / Dice counts, constant 4.19 Mpixel per
          batch at 256^2 / 512^2 / 1024^2; "replicas only" (every rank scores its own batches).
Data parallelism (N > 1): batch shards, sync-BN statistics and gradient all-reduce (parallel.allreduce_gradients: the large
generator gradients in place, asynchronously, back to back; per-rank-mean losses -> grad_scale = 1 / world).
"""
import torch


def _lung_masks(batch):
    yy, xx = torch.meshgrid(torch.arange(256), torch.arange(256), indexing="ij")
    lungs = ((((xx - 85) / 50.0) ** 2 + ((yy - 128) / 80.0) ** 2 < 1) | (((xx - 171) / 45.0) ** 2 + ((yy - 130) / 85.0) ** 2 < 1))
    return lungs.float()[None, None].repeat(batch, 1, 1, 1).contiguous()


def _gan_models(dev, world, capturable=False):
    import semantic_segmentation_b200 as S
    from semantic_segmentation_b200 import gan_program, parallel
    from semantic_segmentation_b200.models_pix2pix import networks
    from semantic_segmentation_b200.optim import Adam
    S.set_precision("bf16")
    gan_program.set_capturable(capturable)      # dropout call count in device memory: the step can be replayed as a CUDA graph
    torch.manual_seed(1234)                      # init_weights draws N(0, 0.02) from torch's generator: same on every rank
    G = networks.define_G(1, 1, 64, "unet_256", "batch", True, "normal", 0.02, [])
    D = networks.define_D(2, 64, "basic", 3, "batch", "normal", 0.02, [])
    G, D = G.to(dev).train(), D.to(dev).train()
    networks.upconv_arch = (1e-3 * torch.randn(8, 3, generator=torch.Generator().manual_seed(5))).to(dev).requires_grad_(True)
    crit = networks.GANLoss("vanilla").to(dev)
    optG = Adam(list(G.parameters()), lr=2e-5, betas=(0.5, 0.999), capturable=capturable)
    optD = Adam(list(D.parameters()), lr=2e-5, betas=(0.5, 0.999), capturable=capturable)
    optG.grad_scale = optD.grad_scale = 1.0 / world      # GANLoss / L1 are means over the rank's shard
    return networks, parallel, G, D, crit, optG, optD


def pix2pix(dev, per, rank, world):
    networks, parallel, G, D, crit, optG, optD = _gan_models(dev, world, capturable=True)
    g = torch.Generator().manual_seed(14 + rank)
    host = (torch.randint(0, 2, (per, 1, 256, 256), generator=g).float().pin_memory(),
            torch.rand(per, 1, 256, 256, generator=g).pin_memory())
    pG, pD = list(G.parameters()), list(D.parameters())

    def step(batch):
        mask, image = batch
        fake = G(mask)
        for p in pD:
            p.requires_grad_(True)
        optD.zero_grad()
        loss_D = 0.5 * (crit(D(torch.cat([mask, fake], 1).detach()), False) + crit(D(torch.cat([mask, image], 1)), True))
        loss_D.backward()
        parallel.allreduce_gradients(pD)
        optD.step()
        for p in pD:
            p.requires_grad_(False)
        optG.zero_grad()
        loss_G = crit(D(torch.cat([mask, fake], 1)), True) + networks.l1_loss(fake, image) * 100.0
        loss_G.backward()
        parallel.allreduce_gradients(pG)
        optG.step()
        return loss_G.detach()

    return dict(step=step, host_batch=host, units=per, graphable=True,
                metric="Pix2Pix optimize_parameters images/sec @256^2 (unet_256 G + 70x70 PatchGAN D, GANLoss + 100*L1, Adam, bf16)",
                workload="Pix2Pix optimize_parameters (models_pix2pix/pix2pix_model.py:132-143): G fwd, D step, G step; "
                         "define_G(1,1,64,'unet_256','batch',dropout) + define_D(2,64,'basic'), %d images per GPU" % per,
                # merged-kernel form ~157 GF per image (234 GF as the reference executes its three ConvT per level)
                flops_per_unit=157e9, flops_note="merged 8x8 up-cell form (reference form: 234 GF per image)",
                weight_bytes_note="G 272.3 M + D 2.76 M fp32 parameters: one step streams >= 2 (pack) + 4 (grad write/read) + "
                                  "28 (Adam) bytes per parameter = 9.4 GB -> 1.4 ms at the measured HBM bandwidth, batch independent")


def genseg(dev, per, rank, world):
    import semantic_segmentation_b200 as S  # noqa: F401
    from semantic_segmentation_b200.optim import RMSprop
    from semantic_segmentation_b200.unet import UNet
    from semantic_segmentation_b200.util import augment
    from semantic_segmentation_b200.util.dice_score import seg_loss
    networks, parallel, G, D, crit, optG, optD = _gan_models(dev, world, capturable=True)
    torch.manual_seed(4321)
    net = UNet(1, 1, False).to(dev).train()          # --classes 1, ConvTranspose2d U-Net (scripts/train_end2end_jsrt.sh)
    pU = list(net.parameters())
    optU = RMSprop(pU, lr=1e-5, weight_decay=1e-8, momentum=0.9)       # seg_loss is a global-batch mean: grad_scale 1
    g = torch.Generator().manual_seed(14 + rank)
    mask = _lung_masks(per)
    # the affine matrices of the mask augmentation are part of the batch (drawn on the host like the images), so the whole
    # iteration reads only device-resident inputs and can be replayed as one CUDA graph
    host = (mask.pin_memory(), torch.rand(per, 1, 256, 256, generator=g).pin_memory(),
            torch.rand(per, 1, 256, 256, generator=g).pin_memory(),
            augment.draw_affine(per, 256, 256, torch.Generator().manual_seed(100 + 8 * rank)).pin_memory())
    pG, pD = list(G.parameters()), list(D.parameters())

    def step(batch):
        mask, image, val_image, mat = batch
        for p in pD:                                   # Generator problem (train_end2end_jsrt.py:141-155)
            p.requires_grad_(False)
        optG.zero_grad()
        fake = G(mask)
        loss_G = crit(D(torch.cat([mask, fake], 1)), True) + networks.l1_loss(fake, image) * 100.0
        loss_G.backward()
        parallel.allreduce_gradients(pG)
        optG.step()
        for p in pD:                                   # Discriminator problem (:158-173)
            p.requires_grad_(True)
        optD.zero_grad()
        with torch.no_grad():
            fake = G(mask)
        loss_D = 0.5 * (crit(D(torch.cat([mask, fake], 1)), False) + crit(D(torch.cat([mask, image], 1)), True))
        loss_D.backward()
        parallel.allreduce_gradients(pD)
        optD.step()
        optU.zero_grad()                               # Unet problem (:177-232): real batch + generated batch
        loss = seg_loss(net(image), mask.long())
        fake_mask = augment.augment_masks(mask, mat)
        with torch.no_grad():
            fake_image = augment.synth_image_postprocess(G(fake_mask))
        fake_loss = seg_loss(net(fake_image), fake_mask.long())
        (loss + 1.0 * fake_loss).backward()
        parallel.allreduce_gradients(pU)
        optU.step()
        vi = val_image.clone().requires_grad_(True)    # Arch problem's training_step on the validation batch
        seg_loss(net(vi), mask.long()).backward()
        return loss_G.detach()

    return dict(step=step, host_batch=host, units=per, graphable=True,
                metric="GenSeg end-to-end iteration, real images/sec @256^2 (G step + D step + U-Net step on real + generated batch "
                       "+ Arch training_step, bf16)",
                workload="GenSeg end-to-end iteration (running_files/train_end2end_jsrt.py:141-236), %d real images per GPU; the "
                         "mask augmentation runs on the device (util/augment.py) in place of the CPU imgaug round trip; the DARTS "
                         "hyper-gradient of the external betty-ml engine is not included" % per,
                flops_per_unit=None, flops_note=None, weight_bytes_note=None)


def evaluate(dev, n, hw):
    import semantic_segmentation_b200 as S
    from semantic_segmentation_b200.unet import UNet
    from semantic_segmentation_b200.unet.evaluate import dice_from_counts, predict_counts
    S.set_precision("bf16")
    torch.manual_seed(1234)
    net = UNet(1, 2, True).to(dev).eval()
    g = torch.Generator().manual_seed(3)
    host = (torch.rand(n, 1, hw, hw, generator=g).pin_memory(), torch.randint(0, 2, (n, hw, hw), generator=g).pin_memory())

    def step(batch):
        x, y = batch
        with torch.no_grad():
            counts, _ = predict_counts(net, x, y)
            return dice_from_counts(counts)

    return dict(step=step, host_batch=host, units=n, graphable=True,
                metric="U-Net inference images/sec (eval forward + fused argmax / Dice counts, bf16)",
                workload="UNet(1,2,bilinear) evaluate path (unet/evaluate.py:10-43): %d x 1 x %d x %d per batch" % (n, hw, hw),
                flops_per_unit=79.818e9 * (hw / 256.0) ** 2, flops_note="forward only", weight_bytes_note=None)
