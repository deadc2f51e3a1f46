"""NVMe tier for optimizer state (ZeRO-Infinity): fp32 master / exp_avg / exp_avg_sq of a sub-group live in files and are
streamed through pinned host buffers around that sub-group's step, reads of sub-group i+1 overlapping the step of i.
Parity: deepspeed/runtime/swap_tensor/partitioned_optimizer_swapper.py:29, pipelined_optimizer_swapper.py:51."""
import os

import torch

from ...ops.aio import aio_handle


class OptimizerStateSwapper:

    def __init__(self, swap_folder, aio_config, numel_per_group, dtype=torch.float32, rank=0, state_names=("master", "exp_avg", "exp_avg_sq")):
        self.folder = os.path.join(str(swap_folder), "optimizer", f"rank{rank}")
        os.makedirs(self.folder, exist_ok=True)
        self.handle = aio_handle(aio_config.get("block_size", 1 << 20), aio_config.get("queue_depth", 8),
                                 aio_config.get("single_submit", False), aio_config.get("overlap_events", True),
                                 aio_config.get("thread_count", 1))
        self.read_handle = aio_handle(aio_config.get("block_size", 1 << 20), aio_config.get("queue_depth", 8),
                                      aio_config.get("single_submit", False), aio_config.get("overlap_events", True),
                                      aio_config.get("thread_count", 1))
        self.numel_per_group = list(numel_per_group)
        self.state_names = tuple(state_names)
        self.dtype = dtype
        n = max(self.numel_per_group)
        pin = torch.cuda.is_available()
        mk = lambda: (torch.empty(n, dtype=dtype).pin_memory() if pin else torch.empty(n, dtype=dtype))
        self.buffers = [{s: mk() for s in self.state_names} for _ in range(2)]  # double buffered
        self._prefetched = None

    def _path(self, group, name):
        return os.path.join(self.folder, f"group{group}_{name}.swp")

    def initialize(self, group, tensors):
        """Write the initial state of `group` (dict name -> host tensor)."""
        for name, t in tensors.items():
            self.handle.sync_pwrite(t.contiguous(), self._path(group, name))

    def prefetch(self, group):
        buf = self.buffers[group % 2]
        n = self.numel_per_group[group]
        for name in self.state_names:
            self.read_handle.async_pread(buf[name][:n], self._path(group, name))
        self._prefetched = group

    def swap_in(self, group):
        if self._prefetched != group:
            self.prefetch(group)
        self.read_handle.wait()
        self._prefetched = None
        n = self.numel_per_group[group]
        return {name: self.buffers[group % 2][name][:n] for name in self.state_names}

    def swap_out(self, group, async_op=True):
        n = self.numel_per_group[group]
        for name in self.state_names:
            self.handle.async_pwrite(self.buffers[group % 2][name][:n], self._path(group, name))
        if not async_op:
            self.handle.wait()

    def flush(self):
        self.handle.wait()
