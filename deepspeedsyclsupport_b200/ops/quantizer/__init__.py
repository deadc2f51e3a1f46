from .quantizer import *  # noqa: F401,F403
from .quantizer import Quantizer, ds_quantizer  # noqa: F401
