"""Bookkeeping shared by the optimizer-state swappers (parity: deepspeed/runtime/swap_tensor/optimizer_utils.py:21-476): where a
parameter's state tensors and gradient partitions live on disk, and the base class that owns the AIO configuration, alignment
arithmetic and the gradient swap-out path.  The streaming of whole sub-groups around the optimizer step is
optimizer_swapper.OptimizerStateSwapper."""
import os

import torch

from ...utils.logging import logger
from .utils import AIO_ALIGNED_BYTES, MIN_AIO_BYTES, swap_in_tensors


class FlattenedTensorSwapInfo(object):
    """A slice [offset, offset + length) of a parameter's flat gradient, stored at `path`."""

    def __init__(self, path, length, offset):
        self.path = path
        self.offset = offset
        self.length = length


class OptimizerStateSwapInfo(object):

    def __init__(self, parameter, numel, base_folder):
        self.tensors = []
        self.param_id = OptimizerSwapper.parameter_id(parameter)
        self.swap_folder = base_folder
        self.swap_paths = []
        self.swapped_gradients = {}
        self.unswapped_gradients = {}
        self.tensor_numel = numel
        self.tensor_dtype = parameter.dtype
        self.tensor_device = parameter.device
        self.has_state_tensors = False
        self._add_tensors([parameter])

    def numel(self):
        return self.tensor_numel

    def has_gradients(self):
        return bool(self.swapped_gradients) or bool(self.unswapped_gradients)

    def _add_tensors(self, tensor_list):
        for t in tensor_list:
            self.tensors.append(t)
            self.swap_paths.append(os.path.join(self.swap_folder, f"{self.param_id}_tensor_{len(self.swap_paths)}.tensor.swp"))

    def add_state_tensors(self, tensor_list):
        self.has_state_tensors = True
        self._add_tensors(tensor_list)

    def device(self):
        return self.tensor_device

    def dtype(self):
        return self.tensor_dtype

    def release_memory(self):
        for t in self.tensors:
            t.data = torch.Tensor()

    def get_or_create_gradient_paths(self, offsets, lengths):
        paths = []
        for offset, length in zip(offsets, lengths):
            if offset not in self.swapped_gradients:
                path = os.path.join(self.swap_folder, f"{self.param_id}_gradient_{offset}_{length}.tensor.swp")
                self.swapped_gradients[offset] = FlattenedTensorSwapInfo(path, length, offset)
            paths.append(self.swapped_gradients[offset].path)
        return paths

    def set_swap_buffers(self, buffers):
        compute_lengths = [self.numel()] * len(self.tensors)
        for t, buf, n in zip(self.tensors, buffers, compute_lengths):
            t.data = buf.data.narrow(0, 0, n)

    def get_swap_gradient_buffers(self, swap_buffer):
        assert self.numel() <= swap_buffer.numel()
        return [swap_buffer.narrow(0, g.offset, g.length) for g in self.swapped_gradients.values()]

    def get_swap_gradient_paths(self):
        return [g.path for g in self.swapped_gradients.values()]

    def get_unpinned_state_tensors(self):
        return [t for t in self.tensors[1:] if not t.is_pinned()]

    def read_unswapped_gradients(self, dest_buffer):
        num_elem = 0
        for offset, grad in self.unswapped_gradients.items():
            dest_buffer.narrow(0, offset, grad.numel()).copy_(grad.reshape(-1))
            num_elem += grad.numel()
        return num_elem

    def release_unswapped_gradients(self):
        self.unswapped_gradients = {}


SWAPPER_DEBUG_MODE = False
SWAP_OUT_GRADIENT_TIMER = "swap_out_gradient"


class OptimizerSwapper(object):

    @staticmethod
    def parameter_id(param):
        return getattr(param, "ds_id", id(param))

    def __init__(self, swap_config, aio_config, base_folder, optimizer, largest_numel, device, dtype, timers=None):
        self.swap_config = swap_config
        self.aio_config = aio_config
        self.swap_params_info = {}
        self.swap_element_size = torch.tensor([], dtype=dtype).element_size()
        self.swap_folder = os.path.join(base_folder, "optimizer", f"rank{int(os.environ.get('RANK', 0))}")
        os.makedirs(self.swap_folder, exist_ok=True)
        self.optimizer = optimizer
        block = aio_config.get("block_size", 1 << 20) if isinstance(aio_config, dict) else aio_config.block_size
        threads = aio_config.get("thread_count", 1) if isinstance(aio_config, dict) else aio_config.thread_count
        self.min_aio_bytes = max(MIN_AIO_BYTES, block)
        self.aligned_bytes = AIO_ALIGNED_BYTES * threads
        self.numel_alignment = self.aligned_bytes // self.swap_element_size
        self.largest_numel = self._io_aligned_numel(largest_numel)
        self.dtype = dtype
        self.device = device
        self.timers = timers
        self.timer_names = set()
        self.print_exclude_list = ["optimizer", "swap_buffer_manager", "swap_params_info", "timers", "timer_names"]

    def swappable_tensor(self, param=None, numel=None):
        assert param is not None or numel is not None, "Either param or numel must be provided"
        n = param.numel() if param is not None else numel
        return self.min_aio_bytes <= n * self.swap_element_size

    def init_timers(self):
        self.timer_names = set()

    def log_timers(self):
        if self.timer_names and self.timers:
            self.timers.log(list(self.timer_names))

    def pre_backward(self):
        self.init_timers()

    def post_backward(self):
        pass

    def _io_aligned_numel(self, numel):
        r = numel % self.numel_alignment
        return numel if r == 0 else numel + self.numel_alignment - r

    def _get_param_swap_info(self, parameter):
        return self.swap_params_info.get(self.parameter_id(parameter))

    def _create_param_swap_info(self, parameter, numel):
        pid = self.parameter_id(parameter)
        assert pid not in self.swap_params_info
        info = self.swap_params_info[pid] = OptimizerStateSwapInfo(parameter, numel, self.swap_folder)
        state = self._get_state_tensors(parameter)
        if state:
            info.add_state_tensors(state)
        return info

    def _get_state_tensors(self, parameter):
        if self.optimizer is None or parameter not in self.optimizer.state:
            return []
        return [v for v in self.optimizer.state[parameter].values() if torch.is_tensor(v) and v.numel() > 1]

    def _update_param_state_info(self, swap_info, parameter):
        if not swap_info.has_state_tensors:
            state = self._get_state_tensors(parameter)
            if state:
                swap_info.add_state_tensors(state)

    def _swap_out_gradients(self, parameter, gradient_offsets, gradient_tensors, gradient_swapper):
        """Large, aligned gradient slices go to disk through `gradient_swapper` (an AsyncTensorSwapper); the rest stay in memory
        until the optimizer step reads them (`read_unswapped_gradients`)."""
        info = self._get_param_swap_info(parameter)
        if info is None:
            return
        swappable_t, swappable_o, swappable_l = [], [], []
        for offset, t in zip(gradient_offsets, gradient_tensors):
            if not self.swappable_tensor(numel=t.numel()) or t.numel() % self.numel_alignment:
                info.unswapped_gradients[offset] = t
                continue
            swappable_t.append(t)
            swappable_o.append(offset)
            swappable_l.append(t.numel())
        if swappable_t:
            paths = info.get_or_create_gradient_paths(swappable_o, swappable_l)
            gradient_swapper.swap_out_tensors(tensor_list=swappable_t, path_list=paths)
            self.timer_names.add(SWAP_OUT_GRADIENT_TIMER)

    def _flush_gradient_swapper(self, gradient_swapper):
        if gradient_swapper.has_buffers():
            buffers = gradient_swapper.release_buffers()
            self.timer_names.update(gradient_swapper.get_timer_names())
            return buffers
        return []

    def _retrieve_unswapped_grad_partitions(self, swap_info, dest_buffer):
        n = swap_info.read_unswapped_gradients(dest_buffer)
        swap_info.release_unswapped_gradients()
        return n

    def _swap_in_gradients(self, aio_handle, parameter, dest_buffer):
        """Assemble the full flat gradient of `parameter` in `dest_buffer`: swapped slices from disk, the rest from memory."""
        info = self._get_param_swap_info(parameter)
        if info is None or not info.has_gradients():
            return 0
        bufs, paths = info.get_swap_gradient_buffers(dest_buffer), info.get_swap_gradient_paths()
        if bufs:
            swap_in_tensors(aio_handle, bufs, paths)
            aio_handle.wait()
        n = sum(b.numel() for b in bufs) + self._retrieve_unswapped_grad_partitions(info, dest_buffer)
        if SWAPPER_DEBUG_MODE:
            logger.info(f"swapped in {n} gradient elements of parameter {info.param_id}")
        return n
