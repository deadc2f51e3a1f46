"""Per-interface registries of implementations (parity: inference/v2/modules/module_registry.py:14-58)."""
from abc import ABC, abstractmethod
from typing import Any, Dict, Type

from ....runtime.config_utils import DeepSpeedConfigModel
from .ds_module import DSModuleBase


class ConfigBundle(DeepSpeedConfigModel):
    """Everything needed to pick and build one implementation: its registry key, the interface config, and options that only
    this implementation understands."""
    name: str
    config: DeepSpeedConfigModel
    implementation_config: Dict[str, Any] = {}


class DSModuleRegistryBase(ABC):
    """Subclasses declare `registry: Dict = {}` and `associated_class()`; implementations join with
    `@XRegistry.register_module`."""

    @classmethod
    def instantiate_config(cls, config_bundle: ConfigBundle) -> DSModuleBase:
        if config_bundle.name not in cls.registry:
            raise KeyError(f"Unknown DSModule: {config_bundle.name}, cls.registry={cls.registry}")
        impl = cls.registry[config_bundle.name]
        if not impl.supports_config(config_bundle.config):
            raise ValueError(f"Config {config_bundle.config} is not supported by {impl}")
        return impl(config_bundle.config, config_bundle.implementation_config)

    @staticmethod
    @abstractmethod
    def associated_class() -> Type[DSModuleBase]:
        """Interface every member of this registry implements."""

    @classmethod
    def register_module(cls, child_class):
        if not (isinstance(child_class, type) and issubclass(child_class, cls.associated_class())):
            raise TypeError(f"Can only register subclasses of {cls.associated_class()}, {child_class} does not inherit from it")
        cls.registry[child_class.name()] = child_class
        return child_class
