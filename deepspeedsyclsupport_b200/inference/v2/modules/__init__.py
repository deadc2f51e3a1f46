"""FastGen building blocks: interface per block kind, registry of implementations, heuristics that pick one
(parity: inference/v2/modules/)."""
from . import heuristics, implementations, interfaces
from .configs import (DSEmbeddingsConfig, DSLinearConfig, DSMoEConfig, DSNormConfig, DSSelfAttentionConfig, DSUnembedConfig)
from .module_registry import ConfigBundle, DSModuleRegistryBase
