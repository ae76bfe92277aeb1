"""Parts of the U-Net -- drop-in for the reference's unet/unet_parts.py (same class names, constructor
signatures, sub-module names and state_dict layout), running on the sm_100a kernels.

The torch.nn layers below are PARAMETER CONTAINERS only (they give the reference's state_dict keys,
shapes and default initialisation); their forward() is never called.  The arithmetic runs in
libgenseg_b200.so through ``semantic_segmentation_b200.engine``.
"""
import torch
import torch.nn as nn

from ..functions import run_program
from ..part_programs import DoubleConvProgram, DownProgram, OutConvProgram, UpProgram


def _conv_bn_relu_slots(cin, cout):
    """Three container slots (indices k, k+1, k+2 of ``double_conv``): bias-free 3x3 weight, BatchNorm state, no state."""
    return [nn.Conv2d(cin, cout, 3, padding=1, bias=False), nn.BatchNorm2d(cout), nn.ReLU(True)]


class DoubleConv(nn.Module):
    """Two fused conv3x3 + BatchNorm + ReLU stages (reference: unet/unet_parts.py:8-25); state_dict keys
    ``double_conv.{0,3}.weight`` and ``double_conv.{1,4}.*``."""

    def __init__(self, in_channels, out_channels, mid_channels=None):
        super().__init__()
        mid = mid_channels or out_channels
        self.double_conv = nn.Sequential(*_conv_bn_relu_slots(in_channels, mid), *_conv_bn_relu_slots(mid, out_channels))
        self._program = DoubleConvProgram(in_channels, out_channels, mid)

    def forward(self, x):
        return run_program(self._program, self, x)


class Down(nn.Module):
    """2x2 max-pool feeding a DoubleConv (reference: unet/unet_parts.py:28-39); keys ``maxpool_conv.1.double_conv.*``."""

    def __init__(self, in_channels, out_channels):
        super().__init__()
        self.maxpool_conv = nn.Sequential(nn.MaxPool2d(2), DoubleConv(in_channels, out_channels))
        self._program = DownProgram(in_channels, out_channels)

    def forward(self, x):
        return run_program(self._program, self, x)


class Up(nn.Module):
    """x2 upsample (bilinear, or a k2 s2 transposed convolution with weights ``up.weight / up.bias``) written next to the
    skip in one concat buffer, then a DoubleConv (reference: unet/unet_parts.py:42-68)."""

    def __init__(self, in_channels, out_channels, bilinear=True):
        super().__init__()
        half = in_channels // 2
        self.up = (nn.Upsample(scale_factor=2, mode='bilinear', align_corners=True) if bilinear
                   else nn.ConvTranspose2d(in_channels, half, kernel_size=2, stride=2))
        self.conv = DoubleConv(in_channels, out_channels, half if bilinear else None)
        self._program = UpProgram(in_channels, out_channels, bilinear)

    def forward(self, x1, x2):
        return run_program(self._program, self, (x1, x2))


class OutConv(nn.Module):
    """Biased 1x1 convolution to the class logits, fp32 NCHW out (reference: unet/unet_parts.py:71-77)."""

    def __init__(self, in_channels, out_channels):
        super().__init__()
        self.conv = nn.Conv2d(in_channels, out_channels, kernel_size=1)
        self._program = OutConvProgram(in_channels, out_channels)

    def forward(self, x):
        return run_program(self._program, self, x)
