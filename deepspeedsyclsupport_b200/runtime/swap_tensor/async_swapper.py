"""Asynchronous swap-out of many small tensors through a few large pinned staging buffers (parity:
deepspeed/runtime/swap_tensor/async_swapper.py:19-176).

Typical use: gradient partitions leaving for NVMe while backward keeps producing more.  Tensors are packed back to back (each
start aligned for O_DIRECT) into the current staging buffer; a full buffer is handed to the AIO engine -- one write per tensor,
from its slice of the buffer to its own file -- and packing continues in the next buffer.  Only when every buffer is in flight
does the caller wait."""
import torch

from ...utils.logging import logger
from .utils import MIN_AIO_BYTES, swap_out_tensors

INVALID_BUFFER_INDEX = -1
ASYNC_SWAPPER_WAIT_TIMER = "async_swap_gradient_wait"


class _Staging:
    """One pinned buffer with an append cursor and the (slice, path) pairs packed into it."""

    def __init__(self, buffer):
        self.buffer = buffer
        self.reset()

    def reset(self):
        self.offset, self.tensors, self.paths, self.num_elem = 0, [], [], 0

    def has_space(self, numel):
        return self.offset + numel <= self.buffer.numel()

    def insert(self, tensor, path, aligned_numel):
        dst = self.buffer.narrow(0, self.offset, aligned_numel)
        dst.narrow(0, 0, tensor.numel()).copy_(tensor.reshape(-1))
        self.tensors.append(dst)
        self.paths.append(path)
        self.offset += aligned_numel
        self.num_elem += tensor.numel()


class AsyncTensorSwapper(object):

    def __init__(self, aio_handle, numel_alignment, timers=None):
        self.free_buffer_index = []
        self.swapping_buffer_index = []
        self.ready_buffer_index = []
        self.current_buffer_index = INVALID_BUFFER_INDEX
        self.all_buffers = []
        self.aio_handle = aio_handle
        self.numel_alignment = numel_alignment
        self.max_numel = 0
        self.num_pending_swaps = 0
        self.timers = timers
        self.timer_names = set()
        self.num_elements_swapped = 0
        self.dtype = None

    def has_buffers(self):
        return len(self.all_buffers) > 0

    def add_buffers(self, buffer_list):
        assert len(self.all_buffers) == 0 and all(b.device.type == "cpu" for b in buffer_list)
        dtype = buffer_list[0].dtype
        assert all(b.dtype == dtype for b in buffer_list)
        self.dtype = dtype
        self.all_buffers = [_Staging(b) for b in buffer_list]
        self.free_buffer_index = list(range(len(self.all_buffers)))
        self.max_numel = max(b.numel() for b in buffer_list)
        self.timer_names = set()

    def get_timer_names(self):
        return list(self.timer_names)

    def release_buffers(self):
        """Write out whatever is packed, wait for everything in flight and give the pinned buffers back."""
        self._report_statistics("Swapped out[Before flush]")
        self._flush_buffers_until_complete()
        self._report_statistics("Swapped out[After flush]")
        buffers = [s.buffer for s in self.all_buffers]
        self.all_buffers, self.free_buffer_index, self.current_buffer_index = [], [], INVALID_BUFFER_INDEX
        self.num_elements_swapped = 0
        self.dtype = None
        return buffers

    def swap_out_tensors(self, tensor_list, path_list):
        for tensor, path in zip(tensor_list, path_list):
            self._swap_out_tensor(tensor, path)

    # ---- internals
    def _report_statistics(self, message):
        if self.dtype is not None and self.num_elements_swapped:
            gb = self.num_elements_swapped * torch.empty((), dtype=self.dtype).element_size() / 1024**3
            logger.debug(f"{message} num_elems = {self.num_elements_swapped}, {gb:5.2f} GB")

    def _io_aligned_numel(self, numel):
        r = numel % self.numel_alignment
        return numel if r == 0 else numel + self.numel_alignment - r

    def _swap_out_tensor(self, tensor, swap_path):
        assert len(self.all_buffers) > 0, "add_buffers first"
        aligned = self._io_aligned_numel(tensor.numel())
        assert aligned <= self.max_numel, f"tensor of {tensor.numel()} elements does not fit a staging buffer of {self.max_numel}"
        self._make_swap_space(aligned)
        self._get_current_buffer().insert(tensor, swap_path, aligned)

    def _make_swap_space(self, numel):
        if self.current_buffer_index == INVALID_BUFFER_INDEX:
            self._allocate_buffer()
            return
        if not self._get_current_buffer().has_space(numel):
            # the current buffer is full: it becomes ready; start its writes if a free buffer lets packing continue
            if self.free_buffer_index:
                self._flush_ready_buffers()
            else:
                self._flush_buffers_until_complete()
            self._allocate_buffer()

    def _allocate_buffer(self):
        assert self.current_buffer_index == INVALID_BUFFER_INDEX and self.free_buffer_index
        self.current_buffer_index = self.free_buffer_index.pop()
        self._get_current_buffer().reset()

    def _flush_ready_buffers(self):
        if self.current_buffer_index != INVALID_BUFFER_INDEX:
            self.ready_buffer_index.append(self.current_buffer_index)
            self.current_buffer_index = INVALID_BUFFER_INDEX
        self._swap_out_ready_buffers()

    def _flush_buffers_until_complete(self):
        self._flush_ready_buffers()
        assert len(self.ready_buffer_index) == 0
        self._wait_for_swap_complete()
        assert len(self.swapping_buffer_index) == 0 and len(self.free_buffer_index) == len(self.all_buffers)

    def _swap_out_ready_buffers(self):
        for idx in self.ready_buffer_index:
            s = self.all_buffers[idx]
            if s.tensors:
                swap_out_tensors(self.aio_handle, s.tensors, s.paths)
                self.num_pending_swaps += len(s.tensors)
        self.swapping_buffer_index += self.ready_buffer_index
        self.ready_buffer_index = []

    def _wait_for_swap_complete(self):
        self._start_timer(ASYNC_SWAPPER_WAIT_TIMER)
        if self.num_pending_swaps:
            self.aio_handle.wait()
        self._stop_timer(ASYNC_SWAPPER_WAIT_TIMER)
        self.timer_names.add(ASYNC_SWAPPER_WAIT_TIMER)
        self.num_pending_swaps = 0
        for idx in self.swapping_buffer_index:
            self.num_elements_swapped += self.all_buffers[idx].num_elem
            self.all_buffers[idx].reset()
        self.free_buffer_index += self.swapping_buffer_index
        self.swapping_buffer_index = []

    def _get_buffer(self, index):
        assert index != INVALID_BUFFER_INDEX
        return self.all_buffers[index]

    def _get_current_buffer(self):
        return self._get_buffer(self.current_buffer_index)

    def _start_timer(self, name):
        if self.timers:
            self.timers(name).start()

    def _stop_timer(self, name):
        if self.timers:
            self.timers(name).stop()

    def _log_timers(self, name_list, force=False):
        if self.timers and force:
            self.timers.log(name_list)


__all__ = ["AsyncTensorSwapper", "MIN_AIO_BYTES"]
