"""(parity: inference/v2/modules/implementations/pre_norm/cuda_pre_rms.py:20-79; kernels: csrc/norm RMSNorm with fused
residual add)"""
from typing import Any, Dict, Optional, Tuple

import torch

from ......ops import fused_ops as FO
from ....inference_parameter import InferenceParameter
from ....inference_utils import DtypeEnum, NormTypeEnum
from ...configs import DSNormConfig
from ...interfaces import DSPreNormBase, DSPreNormRegistry

_FP = (DtypeEnum.fp16, DtypeEnum.bf16, DtypeEnum.fp32)


def _dtypes_ok(config: DSNormConfig) -> bool:
    return all(DtypeEnum(d) in _FP for d in (config.residual_dtype, config.input_dtype, config.output_dtype)) and \
        config.residual_dtype == config.input_dtype == config.output_dtype


@DSPreNormRegistry.register_module
class DSPreRMSCUDAModule(DSPreNormBase):

    @staticmethod
    def name() -> str:
        return "cuda_pre_rms"

    @staticmethod
    def supports_config(config: DSNormConfig) -> bool:
        return NormTypeEnum(config.type) == NormTypeEnum.RMSNorm and _dtypes_ok(config)

    def transform_param(self, param: torch.Tensor) -> InferenceParameter:
        return InferenceParameter.initialize(param.to(DtypeEnum(self._config.input_dtype).value).contiguous())

    def forward(self, residual: torch.Tensor, hidden_states: Optional[torch.Tensor], gamma: torch.Tensor,
                beta: Optional[torch.Tensor] = None) -> Tuple[torch.Tensor, torch.Tensor]:
        if hidden_states is None:
            return residual, FO.rms_norm(residual, gamma, self._config.eps)
        out, new_residual = FO.rms_norm(hidden_states, gamma, self._config.eps, residual)
        return new_residual, out
