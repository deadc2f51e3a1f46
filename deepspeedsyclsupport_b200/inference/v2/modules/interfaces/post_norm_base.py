"""(parity: inference/v2/modules/interfaces/post_norm_base.py:18-69)"""
from typing import Any, Dict, Optional, Tuple, Type

import torch

from ...inference_parameter import InferenceParameter
from ..configs import DSNormConfig
from ..ds_module import DSModuleBase
from ..module_registry import DSModuleRegistryBase


class DSPostNormBase(DSModuleBase):
    """Post-norm transformer step:  residual = norm(residual + hidden_states).  Returns (residual, residual)."""

    @staticmethod
    def config_class() -> Type[DSNormConfig]:
        return DSNormConfig

    def __init__(self, config: DSNormConfig, implementation_config: Dict[str, Any] = None) -> None:
        super().__init__(config, implementation_config)

    def transform_param(self, param: torch.Tensor) -> InferenceParameter:
        raise NotImplementedError()

    def forward(self, residual: torch.Tensor, hidden_states: torch.Tensor, gamma: torch.Tensor,
                beta: Optional[torch.Tensor] = None) -> Tuple[torch.Tensor, torch.Tensor]:
        raise NotImplementedError()


class DSPostNormRegistry(DSModuleRegistryBase):
    registry: Dict = {}

    @staticmethod
    def associated_class() -> Type[DSModuleBase]:
        return DSPostNormBase
