from .layers import QUANTIZATION_LAYER_MAPPINGS, QuantizedEmbedding, QuantizedLinear  # noqa: F401
from .quantization import _init_group_wise_weight_quantization  # noqa: F401
from .quantization_context import QuantizationContext  # noqa: F401
from .utils import DeQuantizer, Quantizer  # noqa: F401
