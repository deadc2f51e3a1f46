"""(parity: inference/v2/checkpoint/base_engine.py:14-41)"""
from abc import ABC, abstractmethod
from typing import Iterable, Tuple

import torch

MEGATRON = "megatron"
HUGGINGFACE = "huggingface"


class CheckpointEngineBase(ABC):
    """Yields (checkpoint tensor name, tensor) pairs.  Tensors may come in any order and one at a time: the containers finish a
    parameter as soon as its sources have arrived, so a checkpoint is never held in memory as a whole."""

    @abstractmethod
    def parameters(self) -> Iterable[Tuple[str, torch.Tensor]]:
        ...
