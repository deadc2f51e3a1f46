"""Checkpoint engine interface (parity: deepspeed/runtime/checkpoint_engine/checkpoint_engine.py)."""
import os


class CheckpointEngine:

    def __init__(self, config_params=None):
        pass

    def create(self, tag):
        pass

    def makedirs(self, path, exist_ok=False):
        os.makedirs(path, exist_ok=exist_ok)

    def save(self, state_dict, path: str):
        pass

    def load(self, path: str, map_location=None):
        pass

    def commit(self, tag):
        pass
