"""U-Net drop-in package (reference: unet/__init__.py): ``from semantic_segmentation_b200.unet import UNet``."""
from .unet_model import UNet  # noqa: F401
