"""Base classes of the FastGen building blocks (parity: inference/v2/modules/ds_module.py:14-62).

A DSModule is one implementation of one interface (attention, linear, ...).  It owns no weights: the model passes parameters
to `forward`, after giving the implementation the chance to re-lay them out once (`transform_param`: 16-bit copy, fp8
quantisation, expert stacking)."""
from abc import ABC, abstractmethod
from typing import Any, Dict, Type

import torch

from ....runtime.config_utils import DeepSpeedConfigModel


class DSModuleConfig(DeepSpeedConfigModel):
    max_tokens: int


class DSModuleBase(torch.nn.Module, ABC):

    @staticmethod
    @abstractmethod
    def name() -> str:
        """Key of this implementation in its registry."""

    @staticmethod
    @abstractmethod
    def config_class() -> Type[DSModuleConfig]:
        """Config class of the interface."""

    @staticmethod
    @abstractmethod
    def supports_config(config: DSModuleConfig) -> bool:
        """Whether this implementation can serve `config` (dtypes, shapes, options)."""

    def __init__(self, config: DSModuleConfig, implementation_config: Dict[str, Any] = None) -> None:
        super().__init__()
        self._config = config
        self._implementation_config = implementation_config or {}

    @property
    def config(self):
        return self._config
