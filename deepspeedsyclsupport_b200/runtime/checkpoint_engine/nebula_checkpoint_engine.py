"""Nebula (Azure async tiered checkpoint service) engine hook.  The service SDK (`torch_nebula`) is not available
offline; the class keeps the interface and degrades to synchronous torch.save so configs with a `nebula` block run.
Parity: deepspeed/runtime/checkpoint_engine/nebula_checkpoint_engine.py."""
from ...utils.logging import logger
from .torch_checkpoint_engine import TorchCheckpointEngine


class NebulaCheckpointEngine(TorchCheckpointEngine):

    def __init__(self, config_params=None):
        super().__init__(config_params)
        self.config = config_params
        try:
            import torch_nebula  # noqa: F401
            self.nebula = torch_nebula
        except ImportError:
            self.nebula = None
            logger.warning("torch_nebula is not installed; Nebula checkpoint engine falls back to torch.save")
