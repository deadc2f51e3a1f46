"""Policies: everything family-specific about turning a checkpoint into a running model (parity:
inference/v2/model_implementations/inference_policy_base.py:30-220).

A policy knows (1) how to translate the checkpoint's config into an ArchSpec, (2) which container classes describe the family's
tensors, and (3) under which name prefixes those tensors live (`ContainerMap`).  `build_model` instantiates the model, streams
the checkpoint through the map and verifies that nothing is missing."""
import re
from abc import ABC, abstractmethod
from typing import Dict, Iterable, List, Optional

import torch

from ..checkpoint.base_engine import CheckpointEngineBase
from ..config_v2 import RaggedInferenceEngineConfig
from ..logging import inference_logger

POLICIES: Dict[str, type] = {}


class ContainerMap:
    """Routes checkpoint tensor names to containers by prefix:
         transformer prefix  "model.layers"  : "model.layers.<i>.<rest>" -> transformer_containers[i].set_dependency(rest)
         non-transformer     ["model.embed_tokens", "model.norm", "lm_head"] -> one container, full name
         unmapped            names that are known to be unused (rotary inv_freq buffers, masks)"""

    def __init__(self) -> None:
        self._prefix: Optional[str] = None
        self._transformer: Optional[List] = None
        self._non_transformer_prefixes: List[str] = []
        self._non_transformer = None
        self._unmapped: List[str] = []

    def set_transformer_params(self, prefixes, containers: List) -> None:
        if isinstance(prefixes, (list, tuple)):
            assert len(prefixes) == 1, "one transformer prefix per model"
            prefixes = prefixes[0]
        self._prefix, self._transformer = prefixes.rstrip("."), containers
        self._rx = re.compile("^" + re.escape(self._prefix) + r"\.(\d+)\.(.*)$")

    def set_non_transformer_params(self, container) -> None:
        self._non_transformer = container

    def set_unmapped_params(self, names: Iterable[str]) -> None:
        self._unmapped = list(names)

    @property
    def transformer_params(self):
        return self._transformer

    @property
    def non_transformer_params(self):
        return self._non_transformer

    def map_param(self, name: str, tensor: torch.Tensor) -> None:
        if any(name == u or name.endswith("." + u) for u in self._unmapped):
            return
        m = self._rx.match(name) if self._prefix is not None else None
        if m:
            if not self._transformer[int(m.group(1))].set_dependency(m.group(2), tensor):
                raise ValueError(f"Cannot find container for {name}")
            return
        if self._non_transformer is None or not self._non_transformer.set_dependency(name, tensor):
            raise ValueError(f"Cannot find container for {name}")

    def validate(self) -> None:
        for i, c in enumerate(self._transformer or []):
            if not c.is_populated:
                raise ValueError(f"Transformer container {i} not fully initialized: missing {c.missing()}")
        if self._non_transformer is not None and not self._non_transformer.is_populated:
            raise ValueError(f"Non-transformer container not fully initialized: missing {self._non_transformer.missing()}")


class InferenceV2Policy(ABC):
    model_types: tuple = ()  # HF `model_type` values served by this policy

    def __init_subclass__(cls, **kwargs):
        super().__init_subclass__(**kwargs)
        for t in cls.model_types:
            POLICIES[t] = cls

    def __init__(self, model_config: dict, checkpoint_engine: Optional[CheckpointEngineBase] = None, inf_checkpoint_path: Optional[str] = None) -> None:
        self._model_config = model_config if isinstance(model_config, dict) else model_config.to_dict()
        self._checkpoint_engine = checkpoint_engine
        self._inf_checkpoint_path = inf_checkpoint_path
        self.model = None
        self._dtype = torch.bfloat16

    @abstractmethod
    def instantiate_model(self, engine_config: RaggedInferenceEngineConfig, mp_group, device=None):
        """The family's model with empty containers."""

    @abstractmethod
    def build_container_map(self) -> ContainerMap:
        """ContainerMap over `self.model`'s containers."""

    def rename(self, name: str) -> str:
        """Hook for checkpoints saved from the bare model (without the CausalLM wrapper's prefix)."""
        return name

    def prepare_tensor(self, name: str, tensor: torch.Tensor) -> torch.Tensor:
        """Hook for layouts that need the whole tensor changed before routing (GPT-2's Conv1D transposes)."""
        return tensor

    def build_model(self, engine_config: RaggedInferenceEngineConfig, mp_group, device=None, dtype=None):
        if dtype is not None:
            self._dtype = dtype
        self.model = self.instantiate_model(engine_config, mp_group, device)
        cmap = self.build_container_map()
        if self._checkpoint_engine is None:
            raise ValueError("a checkpoint engine is required to populate the model")
        for name, tensor in self._checkpoint_engine.parameters():
            name = self.rename(name)
            cmap.map_param(name, self.prepare_tensor(name, tensor))
        self.finish_loading(cmap)
        cmap.validate()
        inference_logger().debug(f"{type(self).__name__}: model populated")
        return self.model

    def finish_loading(self, cmap: ContainerMap) -> None:
        """After the stream: tied embeddings feed the unembedding from the embedding table."""


def policy_for(model_type: str):
    from . import FAMILIES  # noqa: F401  (importing the family packages registers their policies)
    if model_type not in POLICIES:
        raise ValueError(f"Unsupported model type {model_type}")
    return POLICIES[model_type]
