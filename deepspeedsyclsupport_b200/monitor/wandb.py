"""Weights & Biases backend (parity: deepspeed/monitor/wandb.py)."""
from .. import comm as dist
from ..utils.logging import logger
from .monitor import Monitor


class WandbMonitor(Monitor):

    def __init__(self, wandb_config):
        super().__init__(wandb_config)
        self.enabled = wandb_config.enabled
        if self.enabled and dist.get_rank() == 0:
            try:
                import wandb
                wandb.init(project=wandb_config.project, group=wandb_config.group, entity=wandb_config.team)
                self._wandb = wandb
            except ImportError:
                logger.warning("wandb is not installed: W&B monitor disabled")
                self.enabled = False

    def log(self, data, step=None, commit=None, sync=None):
        if self.enabled and dist.get_rank() == 0:
            return self._wandb.log(data, step=step, commit=commit, sync=sync)

    def write_events(self, event_list):
        if self.enabled and dist.get_rank() == 0:
            for event in event_list:
                label, value, step = event[0], event[1], event[2]
                self.log({label: value}, step=step)
