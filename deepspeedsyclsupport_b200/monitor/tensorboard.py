"""TensorBoard backend (parity: deepspeed/monitor/tensorboard.py)."""
import os

from .. import comm as dist
from ..utils.logging import logger
from .monitor import Monitor


class TensorBoardMonitor(Monitor):

    def __init__(self, tensorboard_config):
        super().__init__(tensorboard_config)
        self.enabled = tensorboard_config.enabled
        self.output_path = tensorboard_config.output_path
        self.job_name = tensorboard_config.job_name
        self.summary_writer = None
        if self.enabled and dist.get_rank() == 0:
            self.get_summary_writer()

    def get_summary_writer(self, base=os.path.join(os.path.expanduser("~"), "tensorboard")):
        try:
            from torch.utils.tensorboard import SummaryWriter
        except ImportError:
            logger.warning("tensorboard is not installed: TensorBoard monitor disabled")
            self.enabled = False
            return None
        log_dir = os.path.join(self.output_path, self.job_name) if self.output_path else os.path.join(base, self.job_name)
        os.makedirs(log_dir, exist_ok=True)
        self.summary_writer = SummaryWriter(log_dir=log_dir)
        return self.summary_writer

    def write_events(self, event_list, flush=True):
        if self.enabled and self.summary_writer is not None and dist.get_rank() == 0:
            for event in event_list:
                self.summary_writer.add_scalar(*event)
            if flush:
                self.summary_writer.flush()

    def flush(self):
        if self.summary_writer is not None:
            self.summary_writer.flush()
