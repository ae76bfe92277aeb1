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


class DoubleConv(nn.Module):
    """(convolution => [BN] => ReLU) * 2   (reference: unet/unet_parts.py:8-25)"""

    def __init__(self, in_channels, out_channels, mid_channels=None):
        super().__init__()
        if not mid_channels:
            mid_channels = out_channels
        self.double_conv = nn.Sequential(
            nn.Conv2d(in_channels, mid_channels, kernel_size=3, padding=1, bias=False),
            nn.BatchNorm2d(mid_channels),
            nn.ReLU(inplace=True),
            nn.Conv2d(mid_channels, out_channels, kernel_size=3, padding=1, bias=False),
            nn.BatchNorm2d(out_channels),
            nn.ReLU(inplace=True),
        )
        self._program = DoubleConvProgram(in_channels, out_channels, mid_channels)

    def forward(self, x):
        return run_program(self._program, self, x)


class Down(nn.Module):
    """Downscaling with maxpool then double conv   (reference: unet/unet_parts.py:28-39)"""

    def __init__(self, in_channels, out_channels):
        super().__init__()
        self.maxpool_conv = nn.Sequential(nn.MaxPool2d(2), DoubleConv(in_channels, out_channels))
        self._program = DownProgram(in_channels, out_channels)

    def forward(self, x):
        return run_program(self._program, self, x)


class Up(nn.Module):
    """Upscaling then double conv   (reference: unet/unet_parts.py:42-68)"""

    def __init__(self, in_channels, out_channels, bilinear=True):
        super().__init__()
        if bilinear:
            self.up = nn.Upsample(scale_factor=2, mode='bilinear', align_corners=True)
            self.conv = DoubleConv(in_channels, out_channels, in_channels // 2)
        else:
            self.up = nn.ConvTranspose2d(in_channels, in_channels // 2, kernel_size=2, stride=2)
            self.conv = DoubleConv(in_channels, out_channels)
        self._program = UpProgram(in_channels, out_channels, bilinear)

    def forward(self, x1, x2):
        return run_program(self._program, self, (x1, x2))


class OutConv(nn.Module):
    """1x1 conv with bias   (reference: unet/unet_parts.py:71-77)"""

    def __init__(self, in_channels, out_channels):
        super(OutConv, self).__init__()
        self.conv = nn.Conv2d(in_channels, out_channels, kernel_size=1)
        self._program = OutConvProgram(in_channels, out_channels)

    def forward(self, x):
        return run_program(self._program, self, x)
