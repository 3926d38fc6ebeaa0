from .unet import UNet, count_parameters  # noqa: F401
