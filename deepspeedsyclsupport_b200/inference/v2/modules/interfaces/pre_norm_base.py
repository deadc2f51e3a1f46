"""(parity: inference/v2/modules/interfaces/pre_norm_base.py:18-73)"""
from typing import Any, Dict, Optional, Tuple, Type

import torch

from ...inference_parameter import InferenceParameter
from ..configs import DSNormConfig
from ..ds_module import DSModuleBase
from ..module_registry import DSModuleRegistryBase


class DSPreNormBase(DSModuleBase):
    """Pre-norm transformer step:  residual += hidden_states (skipped when hidden_states is None);  out = norm(residual).
    Returns (residual, out)."""

    @staticmethod
    def config_class() -> Type[DSNormConfig]:
        return DSNormConfig

    def __init__(self, config: DSNormConfig, implementation_config: Dict[str, Any] = None):
        super().__init__(config, implementation_config)

    def transform_param(self, param: torch.Tensor) -> InferenceParameter:
        raise NotImplementedError()

    def forward(self, residual: torch.Tensor, hidden_states: Optional[torch.Tensor], gamma: torch.Tensor,
                beta: Optional[torch.Tensor] = None) -> Tuple[torch.Tensor, torch.Tensor]:
        raise NotImplementedError()


class DSPreNormRegistry(DSModuleRegistryBase):
    registry: Dict = {}

    @staticmethod
    def associated_class() -> Type[DSModuleBase]:
        return DSPreNormBase
