"""(parity: inference/v2/modules/implementations/post_norm/cuda_post_ln.py:19-56)"""
from typing import Optional, Tuple

import torch

from ......ops import fused_ops as FO
from ....inference_parameter import InferenceParameter
from ....inference_utils import DtypeEnum, NormTypeEnum
from ...configs import DSNormConfig
from ...interfaces import DSPostNormBase, DSPostNormRegistry
from ..pre_norm.cuda_pre_rms import _dtypes_ok


@DSPostNormRegistry.register_module
class DSPostLNCUDAModule(DSPostNormBase):

    @staticmethod
    def name() -> str:
        return "cuda_post_ln"

    @staticmethod
    def supports_config(config: DSNormConfig) -> bool:
        return NormTypeEnum(config.type) == NormTypeEnum.LayerNorm and _dtypes_ok(config)

    def transform_param(self, param: torch.Tensor) -> InferenceParameter:
        return InferenceParameter.initialize(param.to(DtypeEnum(self._config.input_dtype).value).contiguous())

    def forward(self, residual: torch.Tensor, hidden_states: torch.Tensor, gamma: torch.Tensor,
                beta: Optional[torch.Tensor] = None) -> Tuple[torch.Tensor, torch.Tensor]:
        out, _ = FO.layer_norm(hidden_states, gamma, beta, self._config.eps, residual)
        return out, out
