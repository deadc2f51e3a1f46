"""(parity: inference/v2/modules/interfaces/moe_base.py:18-91)"""
from typing import Any, Dict, Optional, Type

import torch

from ...inference_parameter import InferenceParameter
from ..configs import DSMoEConfig
from ..ds_module import DSModuleBase
from ..module_registry import DSModuleRegistryBase


class DSMoEBase(DSModuleBase):
    """Gate -> top-k routing -> expert MLPs -> weighted combine, for the tokens of one ragged batch."""

    @staticmethod
    def config_class() -> Type[DSMoEConfig]:
        return DSMoEConfig

    def __init__(self, config: DSMoEConfig, implementation_config: Dict[str, Any] = None) -> None:
        super().__init__(config, implementation_config)

    def transform_gate_param(self, param: torch.Tensor) -> InferenceParameter:
        raise NotImplementedError()

    def transform_moe_mlp_1_param(self, param: torch.Tensor) -> InferenceParameter:
        """param: [n_experts, intermediate (x2 when gated), model_dim]"""
        raise NotImplementedError()

    def transform_moe_mlp_2_param(self, param: torch.Tensor) -> InferenceParameter:
        """param: [n_experts, model_dim, intermediate]"""
        raise NotImplementedError()

    def forward(self, hidden_states: torch.Tensor, gate_w: torch.Tensor, mlp_1_w: torch.Tensor, mlp_2_w: torch.Tensor,
                mlp_1_b: Optional[torch.Tensor] = None, mlp_2_b: Optional[torch.Tensor] = None) -> torch.Tensor:
        raise NotImplementedError()

    @property
    def output(self) -> torch.Tensor:
        raise NotImplementedError()


class DSMoERegistry(DSModuleRegistryBase):
    registry: Dict = {}

    @staticmethod
    def associated_class() -> Type[DSModuleBase]:
        return DSMoEBase
