"""b200seg -- Blackwell-native data-parallel U-Net segmentation trainer.

Same capabilities as the reference's single-script trainer
(``/root/reference/Vaihingen PyTorch 2 (кластер).py``), none of its design.  See DESIGN.md.
"""
from .config import TrainConfig, parse_cli  # noqa: F401
from .models.unet import UNet  # noqa: F401

__version__ = "0.1.0"
__all__ = ["TrainConfig", "parse_cli", "UNet"]
