"""Local HuggingFace checkpoint reader (parity: inference/v2/checkpoint/huggingface_engine.py) -- no network: the path
must be a directory with config.json and *.safetensors / pytorch_model*.bin files."""
import glob
import json
import os

import torch

from .base_engine import CheckpointEngineBase


class HuggingFaceCheckpointEngine(CheckpointEngineBase):

    def __init__(self, model_name_or_path: str, auth_token: str = None):
        assert os.path.isdir(model_name_or_path), f"{model_name_or_path} is not a local directory (no network access)"
        self.path = model_name_or_path
        with open(os.path.join(self.path, "config.json")) as f:
            self.model_config = json.load(f)

    def parameters(self):
        files = sorted(glob.glob(os.path.join(self.path, "*.safetensors")))
        if files:
            from safetensors import safe_open
            for fn in files:
                with safe_open(fn, framework="pt", device="cpu") as f:
                    for k in f.keys():
                        yield k, f.get_tensor(k)
            return
        for fn in sorted(glob.glob(os.path.join(self.path, "pytorch_model*.bin"))):
            sd = torch.load(fn, map_location="cpu", weights_only=True)
            for k, v in sd.items():
                yield k, v
