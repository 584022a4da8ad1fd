from .unet_model import UNet  # noqa: F401  (reference: unet/__init__.py:1)
from .unet_resnet import UNetResNet  # noqa: F401
