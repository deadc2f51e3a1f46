"""Event fan-out to TensorBoard / W&B / CSV on rank 0 (parity: deepspeed/monitor/monitor.py:29-53; backends in tensorboard.py,
wandb.py, csv_monitor.py)."""
import importlib
from abc import ABC, abstractmethod

from .. import comm as dist

# backend attribute on MonitorMaster -> (config section, module, class)
_BACKENDS = {
    "tb_monitor": ("tensorboard", ".tensorboard", "TensorBoardMonitor"),
    "wandb_monitor": ("wandb", ".wandb", "WandbMonitor"),
    "csv_monitor": ("csv_monitor", ".csv_monitor", "csvMonitor"),
}


class Monitor(ABC):

    @abstractmethod
    def __init__(self, monitor_config):
        self.monitor_config = monitor_config

    @abstractmethod
    def write_events(self, event_list):
        pass


def _backend_class(module, cls):
    return getattr(importlib.import_module(module, __package__), cls)


class MonitorMaster(Monitor):
    """Owns one instance of every enabled backend (on rank 0) and forwards each event list to all of them."""

    def __init__(self, monitor_config):
        super().__init__(monitor_config)
        self.enabled = monitor_config.enabled
        on_rank0 = dist.get_rank() == 0
        for attr, (section, module, cls) in _BACKENDS.items():
            cfg = getattr(monitor_config, section)
            setattr(self, attr, _backend_class(module, cls)(cfg) if (on_rank0 and cfg.enabled) else None)

    def write_events(self, event_list):
        if dist.get_rank() != 0:
            return
        for attr in _BACKENDS:
            backend = getattr(self, attr)
            if backend is not None:
                backend.write_events(event_list)


def __getattr__(name):  # the backend classes used to live in this module
    for _, module, cls in _BACKENDS.values():
        if name == cls:
            return _backend_class(module, cls)
    raise AttributeError(name)
