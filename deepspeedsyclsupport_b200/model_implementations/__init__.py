from .diffusers.unet import DSUNet  # noqa: F401
from .diffusers.vae import DSVAE  # noqa: F401
from .transformers.clip_encoder import DSClipEncoder  # noqa: F401
