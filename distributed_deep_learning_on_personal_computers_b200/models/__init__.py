from .unet import UNet, DoubleConv, DownBlock, UpBlock, count_parameters, conv_flops_per_tile  # noqa: F401
