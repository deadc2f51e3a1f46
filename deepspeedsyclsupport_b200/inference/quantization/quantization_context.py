"""`zero.Init` flavour used while a ZeRO-3 partitioned model is being weight-quantized (reference:
deepspeed/inference/quantization/quantization_context.py)."""
from ...runtime.zero import Init


class QuantizationContext(Init):

    def __init__(self, config_dict_or_path, param_swapper=None):
        super().__init__(config_dict_or_path=config_dict_or_path, param_swapper=param_swapper)
