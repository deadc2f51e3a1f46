"""Weight-only quantized drop-ins for nn.Linear / nn.Embedding: int8 / int4 payload + per-group (scale, offset) live as
buffers, the 16-bit weight exists only for the duration of one forward.  On the device the dequantised tile feeds the
tcgen05 GEMM (`ops.gemm`); the memory held per layer is bits/16 of the original.
(reference: deepspeed/inference/quantization/layers.py QuantizedLinear / QuantizedEmbedding)"""
import torch
from torch import nn
from torch.nn import functional as F

from .utils import DeQuantizer, Quantizer

# one quantized copy per original parameter object: tied weights (embedding <-> lm_head) stay tied
quantized_weight_registry = {}


def _quantize_once(weight, config):
    key = (id(weight), config["num_bits"], config["group_size"], config.get("group_dim", 1), config.get("symmetric", False))
    if key not in quantized_weight_registry:
        quantized_weight_registry[key] = Quantizer(config).quantize(weight.data)
    return quantized_weight_registry[key]


class _QuantizedWeight(object):

    def _adopt(self, config, weight):
        self.config = config
        payload, params, shape = _quantize_once(weight, config)
        self.register_buffer("qweight", payload, persistent=True)
        self.register_buffer("qparams", params, persistent=True)
        self.qshape = shape
        self.compute_dtype = weight.dtype
        self.dequantizer = DeQuantizer(config, weight.dtype)

    def dequantized_weight(self):
        return self.dequantizer.dequantize(self.qweight, self.qparams, self.qshape)


class QuantizedLinear(nn.Module, _QuantizedWeight):

    def __init__(self, config, pre_quant_layer: nn.Linear):
        super().__init__()
        self.in_features, self.out_features = pre_quant_layer.in_features, pre_quant_layer.out_features
        self.bias = pre_quant_layer.bias
        self._adopt(config, pre_quant_layer.weight)

    def forward(self, input):
        w = self.dequantized_weight()
        if input.is_cuda and input.dtype == torch.bfloat16 and w.dtype == torch.bfloat16:
            from ...ops.gemm import gemm
            y = gemm(input.reshape(-1, input.shape[-1]), w).view(*input.shape[:-1], self.out_features)
            return y if self.bias is None else y + self.bias
        return F.linear(input, w, self.bias)

    def extra_repr(self):
        return f"in_features={self.in_features}, out_features={self.out_features}, bits={self.config['num_bits']}, group_size={self.config['group_size']}"


class QuantizedEmbedding(nn.Module, _QuantizedWeight):

    def __init__(self, config, pre_quant_layer: nn.Embedding):
        super().__init__()
        assert pre_quant_layer.max_norm is None and not pre_quant_layer.scale_grad_by_freq and not pre_quant_layer.sparse, "Not supported"
        self.num_embeddings, self.embedding_dim = pre_quant_layer.num_embeddings, pre_quant_layer.embedding_dim
        self.padding_idx = pre_quant_layer.padding_idx
        self._adopt(config, pre_quant_layer.weight)

    def forward(self, input):
        return F.embedding(input, self.dequantized_weight(), self.padding_idx)


QUANTIZATION_LAYER_MAPPINGS = {nn.Linear: QuantizedLinear, nn.Embedding: QuantizedEmbedding}
