"""semantic_segmentation_b200 -- B200-native (sm_100a) hot path of GenSeg (importZL/semantic_segmentation):
U-Net forward/backward + CE/BCE+Dice loss, Pix2Pix unet_256 generator + PatchGAN, Dice evaluation.

Public surface mirrors the reference's modules:
    semantic_segmentation_b200.unet.UNet(n_channels, n_classes, bilinear)
    semantic_segmentation_b200.unet.unet_parts.{DoubleConv, Down, Up, OutConv}
    semantic_segmentation_b200.unet.evaluate.evaluate(net, dataloader, device, amp)
    semantic_segmentation_b200.util.dice_score.{dice_coeff, multiclass_dice_coeff, dice_loss, seg_loss}
    semantic_segmentation_b200.models_pix2pix.networks.{define_G, define_D, GANLoss, ...}
"""
from .engine import get_precision, set_precision, set_tensor_core_mode  # noqa: F401

__version__ = "0.1.0"
