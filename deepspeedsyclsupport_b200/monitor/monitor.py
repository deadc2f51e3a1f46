"""Event fan-out to TensorBoard / W&B / CSV on rank 0 (parity: deepspeed/monitor/monitor.py:29-53 and backends)."""
import csv
import os
from abc import ABC, abstractmethod

from .. import comm as dist
from ..utils.logging import logger


class Monitor(ABC):

    @abstractmethod
    def __init__(self, monitor_config):
        self.monitor_config = monitor_config

    @abstractmethod
    def write_events(self, event_list):
        pass


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


class csvMonitor(Monitor):

    def __init__(self, csv_config):
        super().__init__(csv_config)
        self.filenames = []
        self.enabled = csv_config.enabled
        self.output_path = csv_config.output_path
        self.job_name = csv_config.job_name
        self.log_dir = self.setup_log_dir()

    def setup_log_dir(self, base=os.path.join(os.path.expanduser("~"), "csv_monitor")):
        if self.enabled and dist.get_rank() == 0:
            log_dir = os.path.join(self.output_path, self.job_name) if self.output_path else os.path.join(base, self.job_name)
            os.makedirs(log_dir, exist_ok=True)
            return log_dir

    def write_events(self, event_list):
        if self.enabled and dist.get_rank() == 0:
            for event in event_list:
                log_name, value, step = event[0], event[1], event[2]
                header = ["step", log_name.split("/")[-1]]
                fname = os.path.join(self.log_dir, log_name.replace("/", "_") + ".csv")
                new = fname not in self.filenames and not os.path.exists(fname)
                with open(fname, "a+") as f:
                    w = csv.writer(f)
                    if new:
                        w.writerow(header)
                    if fname not in self.filenames:
                        self.filenames.append(fname)
                    w.writerow([step, value])


class MonitorMaster(Monitor):

    def __init__(self, monitor_config):
        super().__init__(monitor_config)
        self.tb_monitor = self.wandb_monitor = self.csv_monitor = None
        self.enabled = monitor_config.enabled
        if dist.get_rank() == 0:
            if monitor_config.tensorboard.enabled:
                self.tb_monitor = TensorBoardMonitor(monitor_config.tensorboard)
            if monitor_config.wandb.enabled:
                self.wandb_monitor = WandbMonitor(monitor_config.wandb)
            if monitor_config.csv_monitor.enabled:
                self.csv_monitor = csvMonitor(monitor_config.csv_monitor)

    def write_events(self, event_list):
        if dist.get_rank() == 0:
            for m in (self.tb_monitor, self.wandb_monitor, self.csv_monitor):
                if m is not None:
                    m.write_events(event_list)
