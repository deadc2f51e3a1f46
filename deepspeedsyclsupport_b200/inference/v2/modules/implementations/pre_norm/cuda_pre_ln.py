"""(parity: inference/v2/modules/implementations/pre_norm/cuda_pre_ln.py:19-69)"""
from typing import Optional, Tuple

import torch

from ......ops import fused_ops as FO
from ....inference_parameter import InferenceParameter
from ....inference_utils import DtypeEnum, NormTypeEnum
from ...configs import DSNormConfig
from ...interfaces import DSPreNormBase, DSPreNormRegistry
from .cuda_pre_rms import _dtypes_ok


@DSPreNormRegistry.register_module
class DSPreLNCUDAModule(DSPreNormBase):

    @staticmethod
    def name() -> str:
        return "cuda_pre_ln"

    @staticmethod
    def supports_config(config: DSNormConfig) -> bool:
        return NormTypeEnum(config.type) == NormTypeEnum.LayerNorm and _dtypes_ok(config)

    def transform_param(self, param: torch.Tensor) -> InferenceParameter:
        return InferenceParameter.initialize(param.to(DtypeEnum(self._config.input_dtype).value).contiguous())

    def forward(self, residual: torch.Tensor, hidden_states: Optional[torch.Tensor], gamma: torch.Tensor,
                beta: Optional[torch.Tensor] = None) -> Tuple[torch.Tensor, torch.Tensor]:
        if hidden_states is None:
            return residual, FO.layer_norm(residual, gamma, beta, self._config.eps)
        out, new_residual = FO.layer_norm(hidden_states, gamma, beta, self._config.eps, residual)
        return new_residual, out
