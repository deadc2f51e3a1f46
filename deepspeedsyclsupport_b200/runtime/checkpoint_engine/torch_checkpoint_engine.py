import torch

from ...utils.logging import log_dist, logger
from .checkpoint_engine import CheckpointEngine


class TorchCheckpointEngine(CheckpointEngine):

    def create(self, tag):
        log_dist(f"[Torch] Checkpoint {tag} is about to be saved!", ranks=[0])

    def save(self, state_dict, path: str):
        torch.save(state_dict, path)

    def load(self, path: str, map_location=None):
        return torch.load(path, map_location=map_location, weights_only=False)

    def commit(self, tag):
        logger.info(f"[Torch] Checkpoint {tag} is ready now!")
        return True
