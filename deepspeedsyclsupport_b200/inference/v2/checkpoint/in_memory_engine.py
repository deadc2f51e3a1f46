"""(parity: inference/v2/checkpoint/in_memory_engine.py:14-40)"""
from typing import Iterable, Tuple

import torch

from .base_engine import CheckpointEngineBase


class InMemoryModelEngine(CheckpointEngineBase):
    """Serves the parameters of a model that already lives in this process (an `nn.Module` or a state dict) -- e.g. the
    hybrid-engine style hand-off from a training model to the FastGen engine, or random-init models in tests."""

    def __init__(self, model) -> None:
        self.model = model
        self.model_config = getattr(getattr(model, "config", None), "to_dict", lambda: None)()

    def parameters(self) -> Iterable[Tuple[str, torch.Tensor]]:
        if isinstance(self.model, dict):
            yield from self.model.items()
        else:
            yield from self.model.state_dict().items()
