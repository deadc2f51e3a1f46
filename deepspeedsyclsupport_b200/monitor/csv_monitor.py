"""CSV backend: one file per event name, rows of (step, value) (parity: deepspeed/monitor/csv_monitor.py)."""
import csv
import os

from .. import comm as dist
from .monitor import Monitor


class csvMonitor(Monitor):

    def __init__(self, csv_config):
        super().__init__(csv_config)
        self.enabled, self.output_path, self.job_name = csv_config.enabled, csv_config.output_path, csv_config.job_name
        self.filenames = []  # files this process has already opened (their header is written)
        self.log_dir = self.setup_log_dir()

    def _writes(self):
        return self.enabled and dist.get_rank() == 0

    def setup_log_dir(self, base=os.path.join(os.path.expanduser("~"), "csv_monitor")):
        if not self._writes():
            return None
        log_dir = os.path.join(self.output_path or base, self.job_name)
        os.makedirs(log_dir, exist_ok=True)
        return log_dir

    def write_events(self, event_list):
        """event = (name, value, step); `a/b/c` is written to a_b_c.csv with the columns (step, c)."""
        if not self._writes():
            return
        for name, value, step, *_ in event_list:
            path = os.path.join(self.log_dir, name.replace("/", "_") + ".csv")
            first_use = path not in self.filenames
            needs_header = first_use and not os.path.exists(path)
            with open(path, "a+") as f:
                out = csv.writer(f)
                if needs_header:
                    out.writerow(["step", name.rsplit("/", 1)[-1]])
                out.writerow([step, value])
            if first_use:
                self.filenames.append(path)
