"""Write-behind of many small tensors to files through a few large pinned staging buffers.

Role of the reference's deepspeed/runtime/swap_tensor/async_swapper.py:19-176 (gradient partitions leaving for NVMe while
backward keeps producing more), built differently: the staging buffers form a pipeline

    idle  --open-->  filling  --seal-->  (writes submitted) in flight  --aio wait-->  idle

kept as three deques of `_Stage` objects.  A tensor is appended to the filling stage at an offset aligned for O_DIRECT; when it
does not fit, the stage is sealed (one AIO write per packed tensor, slice of the pinned buffer -> its own file) and the next
idle stage is opened.  The caller only blocks in `aio_handle.wait()` when no idle stage is left, or when it takes the buffers
back with `release_buffers()`.
"""
from collections import deque

import torch

from ...utils.logging import logger
from .utils import MIN_AIO_BYTES, swap_out_tensors

ASYNC_SWAPPER_WAIT_TIMER = "async_swap_gradient_wait"


class _Stage:
    """A pinned buffer plus what has been packed into it since it was opened."""
    __slots__ = ("buffer", "cursor", "slices", "files", "payload")

    def __init__(self, buffer):
        self.buffer = buffer
        self.open()

    def open(self):
        self.cursor, self.slices, self.files, self.payload = 0, [], [], 0

    def fits(self, span):
        return self.cursor + span <= self.buffer.numel()

    def pack(self, tensor, path, span):
        window = self.buffer.narrow(0, self.cursor, span)
        window[:tensor.numel()].copy_(tensor.reshape(-1))
        self.slices.append(window)
        self.files.append(path)
        self.cursor += span
        self.payload += tensor.numel()


class AsyncTensorSwapper(object):

    def __init__(self, aio_handle, numel_alignment, timers=None):
        self.aio_handle = aio_handle
        self.numel_alignment = numel_alignment
        self.timers = timers
        self._idle, self._inflight, self._filling = deque(), deque(), None
        self._writes_pending = 0
        self._timer_names = set()
        self.num_elements_swapped = 0
        self.max_numel = 0
        self.dtype = None

    # ---- buffer ownership
    def has_buffers(self):
        return self._filling is not None or bool(self._idle) or bool(self._inflight)

    def add_buffers(self, buffer_list):
        assert not self.has_buffers(), "release_buffers() before handing in a new set"
        assert buffer_list and all(b.device.type == "cpu" and b.dtype == buffer_list[0].dtype for b in buffer_list)
        self.dtype = buffer_list[0].dtype
        self.max_numel = max(b.numel() for b in buffer_list)
        self._idle.extend(_Stage(b) for b in buffer_list)
        self._timer_names = set()

    def release_buffers(self):
        """Submit what is packed, wait for every write, and hand the pinned buffers back to the caller."""
        self._log_volume("before flush")
        self._drain()
        self._log_volume("after flush")
        buffers = [s.buffer for s in self._idle]
        self._idle.clear()
        self.num_elements_swapped, self.dtype, self.max_numel = 0, None, 0
        return buffers

    def get_timer_names(self):
        return sorted(self._timer_names)

    # ---- producer side
    def swap_out_tensors(self, tensor_list, path_list):
        assert self.has_buffers(), "add_buffers() first"
        for tensor, path in zip(tensor_list, path_list):
            span = self._io_aligned_numel(tensor.numel())
            assert span <= self.max_numel, f"a tensor of {tensor.numel()} elements does not fit a staging buffer of {self.max_numel}"
            self._stage_for(span).pack(tensor, path, span)

    def _io_aligned_numel(self, numel):
        a = self.numel_alignment
        return (numel + a - 1) // a * a

    def _stage_for(self, span):
        """The filling stage if `span` fits, otherwise seal it and open the next one (waiting for writes only if none is idle)."""
        if self._filling is not None and self._filling.fits(span):
            return self._filling
        self._seal()
        if not self._idle:
            self._wait_writes()
        self._filling = self._idle.popleft()
        self._filling.open()
        return self._filling

    # ---- pipeline transitions
    def _seal(self):
        stage, self._filling = self._filling, None
        if stage is None:
            return
        if stage.slices:
            swap_out_tensors(self.aio_handle, stage.slices, stage.files)
            self._writes_pending += len(stage.slices)
        self._inflight.append(stage)

    def _wait_writes(self):
        self._timed(True)
        if self._writes_pending:
            self.aio_handle.wait()
        self._timed(False)
        self._writes_pending = 0
        while self._inflight:
            stage = self._inflight.popleft()
            self.num_elements_swapped += stage.payload
            stage.open()
            self._idle.append(stage)

    def _drain(self):
        self._seal()
        self._wait_writes()

    # ---- diagnostics
    def _timed(self, start):
        if self.timers:
            t = self.timers(ASYNC_SWAPPER_WAIT_TIMER)
            t.start() if start else t.stop()
            self._timer_names.add(ASYNC_SWAPPER_WAIT_TIMER)

    def _log_volume(self, when):
        if self.dtype is not None and self.num_elements_swapped:
            gib = self.num_elements_swapped * torch.empty((), dtype=self.dtype).element_size() / 2 ** 30
            logger.debug(f"async swap-out [{when}]: {self.num_elements_swapped} elements, {gib:.2f} GiB")


__all__ = ["AsyncTensorSwapper", "MIN_AIO_BYTES"]
