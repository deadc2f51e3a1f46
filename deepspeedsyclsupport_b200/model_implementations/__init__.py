from .diffusers.unet import DSUNet  # noqa: F401
from .diffusers.vae import DSVAE  # noqa: F401
from .transformers.clip_encoder import DSClipEncoder  # noqa: F401
from .transformers.ds_transformer import (DeepSpeedBERTInference, DeepSpeedBloomInference, DeepSpeedGPTInference,  # noqa: F401,E402
                                            DeepSpeedLlama2Inference, DeepSpeedMegatronGPTInference, DeepSpeedOPTInference,
                                            DeepSpeedTransformerInference)
