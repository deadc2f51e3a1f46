"""Storage back-end interface used by engine.save_checkpoint / load_checkpoint.

A back-end sees the life of one checkpoint tag as   create(tag) -> save(obj, path)* -> commit(tag)   and serves reads
with load(path).  The torch back-end writes files synchronously; the nebula / async back-ends may defer durability until
commit().  (reference: deepspeed/runtime/checkpoint_engine/checkpoint_engine.py)"""
import abc
import os
from typing import Any, Optional


class CheckpointEngine(abc.ABC):

    def __init__(self, config_params: Optional[Any] = None):
        self.config_params = config_params

    def create(self, tag: str) -> None:
        """Called once before the first save() of a tag."""

    def makedirs(self, path: str, exist_ok: bool = False) -> None:
        os.makedirs(path, exist_ok=exist_ok)

    @abc.abstractmethod
    def save(self, state_dict: Any, path: str) -> None:
        ...

    @abc.abstractmethod
    def load(self, path: str, map_location: Any = None) -> Any:
        ...

    def commit(self, tag: str) -> bool:
        """All files of `tag` were handed to save(); returns True once they are durable."""
        return True
