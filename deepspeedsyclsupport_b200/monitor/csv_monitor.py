"""CSV backend: one file per event name, rows of (step, value) (parity: deepspeed/monitor/csv_monitor.py)."""
import csv
import os

from .. import comm as dist
from .monitor import Monitor


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
