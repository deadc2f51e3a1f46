"""Dense projections (parity: inference/v2/modules/implementations/linear/blas_fp_linear.py:25-103).

`blas_fp_linear`: 16-bit weights on the tcgen05 GEMM, bias + activation (plain or gated) in one pass over the output.
`fp8_linear`: e4m3 weights with one scale per output feature, activations quantised per token, product on the fp8 tensor-core
GEMM (csrc/gemm/gemm_fp8.cu) -- chosen by the heuristics when the engine's quantization_mode is "wf8af8".
`fp8_block_linear`: block-scaled e4m3 weights and activations (one fp32 scale per 128 input features of every row, the ZeRO++
wire format) on csrc/gemm/gemm_fp8_block.cu -- quantization_mode "wf8af8_block".
`mxfp8_linear`: MXFP8 (e4m3 + one UE8M0 scale per 32 input features) on the hardware block-scaled tensor-core path
(tcgen05.mma.kind::mxf8f6f4.block_scale, csrc/gemm/gemm_mxfp8.cu) -- quantization_mode "wf8af8_mx"."""
from typing import Any, Dict, Optional

import torch

from ......ops import fused_ops as FO
from ......ops.gemm import gemm, linear_fp8, linear_fp8_block, linear_mxfp8, quantize_fp8_rowwise, quantize_mxfp8
from ....inference_parameter import InferenceParameter
from ....inference_utils import ActivationType, DtypeEnum, activation_name, is_gated
from ...configs import DSLinearConfig
from ...interfaces import DSLinearBase, DSLinearRegistry


def _supported(config: DSLinearConfig) -> bool:
    if config.input_dtype != config.output_dtype:
        return False
    if DtypeEnum(config.input_dtype) not in (DtypeEnum.fp16, DtypeEnum.bf16, DtypeEnum.fp32):
        return False
    return ActivationType(config.activation) != ActivationType.InvalidType


class _LinearCommon(DSLinearBase):

    def __init__(self, config: DSLinearConfig, implementation_config: Dict[str, Any] = None) -> None:
        super().__init__(config, implementation_config)
        self._act = ActivationType(config.activation)
        self._gated = is_gated(self._act)
        self._act_name = activation_name(self._act)
        self._output = None

    @property
    def output(self) -> torch.Tensor:
        return self._output

    def _epilogue(self, y, b):
        if self._gated:
            if b is not None:
                y = y + b
            y = FO.gated_act(y, self._act_name)
        elif self._act != ActivationType.IDENTITY:
            y = FO.bias_act(y, b, self._act_name)
        elif b is not None:
            y = y + b
        self._output = y
        return y


@DSLinearRegistry.register_module
class BlasFPLinear(_LinearCommon):

    @staticmethod
    def name() -> str:
        return "blas_fp_linear"

    @staticmethod
    def supports_config(config: DSLinearConfig) -> bool:
        return _supported(config)

    def transform_param(self, param: torch.Tensor) -> InferenceParameter:
        return InferenceParameter.initialize(param.to(DtypeEnum(self._config.input_dtype).value).contiguous())

    def forward(self, hidden_states: torch.Tensor, w: torch.Tensor, b: Optional[torch.Tensor] = None) -> torch.Tensor:
        if hidden_states.is_cuda and hidden_states.dtype == torch.bfloat16:
            y = gemm(hidden_states, w)
        else:
            y = torch.nn.functional.linear(hidden_states, w)
        return self._epilogue(y, b)


@DSLinearRegistry.register_module
class FP8Linear(_LinearCommon):

    @staticmethod
    def name() -> str:
        return "fp8_linear"

    @staticmethod
    def supports_config(config: DSLinearConfig) -> bool:
        return _supported(config) and config.in_channels % 16 == 0

    def transform_param(self, param: torch.Tensor) -> InferenceParameter:
        if param.dim() != 2:  # biases stay in the activation dtype
            return InferenceParameter.initialize(param.to(DtypeEnum(self._config.input_dtype).value).contiguous())
        q, scale = quantize_fp8_rowwise(param.to(DtypeEnum(self._config.input_dtype).value))
        return InferenceParameter.initialize(q, scales=scale)

    def forward(self, hidden_states: torch.Tensor, w: torch.Tensor, b: Optional[torch.Tensor] = None) -> torch.Tensor:
        return self._epilogue(linear_fp8(hidden_states, w, w.scales), b)


@DSLinearRegistry.register_module
class FP8BlockLinear(_LinearCommon):

    @staticmethod
    def name() -> str:
        return "fp8_block_linear"

    @staticmethod
    def supports_config(config: DSLinearConfig) -> bool:
        return _supported(config) and config.in_channels % 128 == 0

    def transform_param(self, param: torch.Tensor) -> InferenceParameter:
        if param.dim() != 2:
            return InferenceParameter.initialize(param.to(DtypeEnum(self._config.input_dtype).value).contiguous())
        from ......ops.quantizer import quantize_fp8_block
        w = param.to(DtypeEnum(self._config.input_dtype).value).contiguous()
        q, scale = quantize_fp8_block(w)
        return InferenceParameter.initialize(q.view(w.shape), scales=scale.view(w.shape[0], w.shape[1] // 128))

    def forward(self, hidden_states: torch.Tensor, w: torch.Tensor, b: Optional[torch.Tensor] = None) -> torch.Tensor:
        return self._epilogue(linear_fp8_block(hidden_states, w, w.scales), b)


@DSLinearRegistry.register_module
class MXFP8Linear(_LinearCommon):

    @staticmethod
    def name() -> str:
        return "mxfp8_linear"

    @staticmethod
    def supports_config(config: DSLinearConfig) -> bool:
        return _supported(config) and config.in_channels % 128 == 0

    def transform_param(self, param: torch.Tensor) -> InferenceParameter:
        if param.dim() != 2:
            return InferenceParameter.initialize(param.to(DtypeEnum(self._config.input_dtype).value).contiguous())
        q, sf = quantize_mxfp8(param.to(DtypeEnum(self._config.input_dtype).value).contiguous())
        return InferenceParameter.initialize(q, scales=sf)

    def forward(self, hidden_states: torch.Tensor, w: torch.Tensor, b: Optional[torch.Tensor] = None) -> torch.Tensor:
        return self._epilogue(linear_mxfp8(hidden_states, w, w.scales), b)
