"""Event fan-out to TensorBoard / W&B / CSV on rank 0 (parity: deepspeed/monitor/monitor.py:29-53; backends in tensorboard.py,
wandb.py, csv_monitor.py)."""
from abc import ABC, abstractmethod

from .. import comm as dist


class Monitor(ABC):

    @abstractmethod
    def __init__(self, monitor_config):
        self.monitor_config = monitor_config

    @abstractmethod
    def write_events(self, event_list):
        pass


class MonitorMaster(Monitor):

    def __init__(self, monitor_config):
        super().__init__(monitor_config)
        self.tb_monitor = self.wandb_monitor = self.csv_monitor = None
        self.enabled = monitor_config.enabled
        if dist.get_rank() == 0:
            from .csv_monitor import csvMonitor
            from .tensorboard import TensorBoardMonitor
            from .wandb import WandbMonitor
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


def __getattr__(name):  # the backends used to live in this module
    if name == "TensorBoardMonitor":
        from .tensorboard import TensorBoardMonitor
        return TensorBoardMonitor
    if name == "WandbMonitor":
        from .wandb import WandbMonitor
        return WandbMonitor
    if name == "csvMonitor":
        from .csv_monitor import csvMonitor
        return csvMonitor
    raise AttributeError(name)
