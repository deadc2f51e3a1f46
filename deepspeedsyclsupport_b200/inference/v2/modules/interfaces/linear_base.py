"""(parity: inference/v2/modules/interfaces/linear_base.py:18-72)"""
from typing import Any, Dict, Optional, Type

import torch

from ...inference_parameter import InferenceParameter
from ..configs import DSLinearConfig
from ..ds_module import DSModuleBase
from ..module_registry import DSModuleRegistryBase


class DSLinearBase(DSModuleBase):
    """y = act(x @ w^T + b); with a gated activation w has 2 * out_channels rows ([gate; up]) and y has out_channels."""

    @staticmethod
    def config_class() -> Type[DSLinearConfig]:
        return DSLinearConfig

    def __init__(self, config: DSLinearConfig, implementation_config: Dict[str, Any] = None) -> None:
        super().__init__(config, implementation_config)

    def transform_param(self, param: torch.Tensor) -> InferenceParameter:
        raise NotImplementedError()

    def forward(self, hidden_states: torch.Tensor, w: torch.Tensor, b: Optional[torch.Tensor] = None) -> torch.Tensor:
        raise NotImplementedError()

    @property
    def output(self) -> torch.Tensor:
        raise NotImplementedError()


class DSLinearRegistry(DSModuleRegistryBase):
    registry: Dict = {}

    @staticmethod
    def associated_class() -> Type[DSModuleBase]:
        return DSLinearBase
