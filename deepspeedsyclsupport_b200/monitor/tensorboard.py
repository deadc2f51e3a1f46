"""TensorBoard backend (parity: deepspeed/monitor/tensorboard.py): scalars written by the global rank 0 only."""
import os

from .. import comm as dist
from ..utils.logging import logger
from .monitor import Monitor

_DEFAULT_ROOT = os.path.join(os.path.expanduser("~"), "tensorboard")


class TensorBoardMonitor(Monitor):

    def __init__(self, tensorboard_config):
        super().__init__(tensorboard_config)
        cfg = tensorboard_config
        self.enabled, self.output_path, self.job_name = cfg.enabled, cfg.output_path, cfg.job_name
        self.summary_writer = None
        if self._active():
            self.get_summary_writer()

    def _active(self):
        return self.enabled and dist.get_rank() == 0

    def get_summary_writer(self, base=_DEFAULT_ROOT):
        """Create the writer under <output_path or ~/tensorboard>/<job_name>; a missing tensorboard package disables the backend."""
        try:
            from torch.utils.tensorboard import SummaryWriter
        except ImportError:
            logger.warning("tensorboard is not installed: TensorBoard monitor disabled")
            self.enabled = False
            return None
        where = os.path.join(self.output_path or base, self.job_name)
        os.makedirs(where, exist_ok=True)
        self.summary_writer = SummaryWriter(log_dir=where)
        return self.summary_writer

    def write_events(self, event_list, flush=True):
        """event = (tag, value, step)"""
        if not self._active() or self.summary_writer is None:
            return
        for tag, value, step, *_ in event_list:
            self.summary_writer.add_scalar(tag, value, step)
        if flush:
            self.flush()

    def flush(self):
        if self.summary_writer is not None:
            self.summary_writer.flush()
