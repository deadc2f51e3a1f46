"""NVMe tier for ZeRO-3 parameter shards: each parameter's 16-bit shard (`param.ds_tensor`) lives in a file between uses and
is staged through a pool of pinned host buffers; reads are issued asynchronously ahead of the fetch and awaited with
`synchronize_reads`, evictions write back asynchronously.
(reference: deepspeed/runtime/swap_tensor/partitioned_param_swapper.py:36 AsyncPartitionedParameterSwapper)"""
import os
from enum import Enum

import torch

from ...ops.aio import aio_handle
from .utils import SwapBufferPool


class PartitionedParamStatus(Enum):
    AVAILABLE = 1      # shard is in memory
    NOT_AVAILABLE = 2  # shard is only in its file
    INFLIGHT = 3       # read issued, not yet awaited


class AsyncPartitionedParameterSwapper(object):

    def __init__(self, ds_config, model_dtype):
        """ds_config: dict-like with `zero_optimization.offload_param` {nvme_path, buffer_count, buffer_size} and `aio`."""
        zero = ds_config["zero_optimization"] if isinstance(ds_config, dict) else ds_config.zero_config
        op = zero["offload_param"] if isinstance(zero, dict) else zero.offload_param
        get = (lambda k, d=None: op.get(k, d)) if isinstance(op, dict) else (lambda k, d=None: getattr(op, k, d))
        aio = (ds_config.get("aio", {}) if isinstance(ds_config, dict) else getattr(ds_config, "aio_config", {})) or {}
        rank = int(os.environ.get("RANK", "0"))
        self.swap_folder = os.path.join(str(get("nvme_path")), "zero_stage_3", f"{str(model_dtype).split('.')[-1]}params", f"rank{rank}")
        os.makedirs(self.swap_folder, exist_ok=True)
        self.dtype = model_dtype
        self.elements_per_buffer = int(get("buffer_size", 100_000_000))
        self.param_buffer_count = int(get("buffer_count", 5))
        mk = lambda: aio_handle(aio.get("block_size", 1 << 20), aio.get("queue_depth", 8), aio.get("single_submit", False),
                                aio.get("overlap_events", True), aio.get("thread_count", 1))
        self.aio_read_handle, self.aio_write_handle = mk(), mk()
        pin = torch.cuda.is_available()
        bufs = [torch.empty(self.elements_per_buffer, dtype=model_dtype) for _ in range(self.param_buffer_count)]
        self.pool = SwapBufferPool([b.pin_memory() if pin else b for b in bufs])
        self.id_to_path, self.id_to_numel = {}, {}
        self.inflight_params, self.available_params = [], set()
        self.pending_writes, self.pending_reads = 0, 0
        self.invalid_buffer = torch.empty(1, dtype=model_dtype)

    # ---- bookkeeping
    def _path(self, param):
        pid = param.ds_id
        if pid not in self.id_to_path:
            self.id_to_path[pid] = os.path.join(self.swap_folder, f"{pid}_param.tensor.swp")
        return self.id_to_path[pid]

    def swappable_tensor(self, param=None, numel=None):
        n = numel if numel is not None else param.ds_tensor.numel()
        return n * torch.empty(0, dtype=self.dtype).element_size() >= 1024  # tiny shards are not worth a file

    def available_swap_in_buffers(self):
        return len(self.pool.free)

    def get_buffer(self, param, numel):
        buf = self.pool.allocate(param.ds_id, numel)
        self.id_to_numel[param.ds_id] = numel
        return buf

    def reserve_available_buffer(self):  # reference name
        return self.available_swap_in_buffers()

    # ---- device/host memory -> file
    def swap_out_and_release(self, params, async_op=False, force_buffer_release=False):
        """Write the shards to their files; the in-memory shard is replaced by a placeholder."""
        for p in params:
            t = p.ds_tensor
            if p.ds_id in self.pool.used and t.data_ptr() == self.pool.buffers[self.pool.used[p.ds_id]].data_ptr():
                src = t
            else:
                src = self.get_buffer(p, t.numel()) if p.ds_id not in self.pool.used else \
                    self.pool.buffers[self.pool.used[p.ds_id]][:t.numel()]
                src.copy_(t.reshape(-1).to("cpu", self.dtype))
            self.id_to_numel[p.ds_id] = t.numel()
            assert self.aio_write_handle.async_pwrite(src, self._path(p)) == 0
            self.pending_writes += 1
        if not async_op:
            self.synchronize_writes()
        for p in params:
            if not async_op or force_buffer_release:
                self._release(p)

    def _release(self, p):
        if p.ds_id in self.pool.used:
            self.pool.release(p.ds_id)
        p.ds_tensor.data = self.invalid_buffer
        p.ds_tensor.status = PartitionedParamStatus.NOT_AVAILABLE
        self.available_params.discard(p.ds_id)

    def synchronize_writes(self):
        if self.pending_writes:
            self.aio_write_handle.wait()
            self.pending_writes = 0

    # ---- file -> pinned buffer
    def swap_in(self, params, async_op=True, swap_in_buffers=None):
        assert len(params) <= self.available_swap_in_buffers() or swap_in_buffers is not None, \
            f"not enough swap buffers ({self.available_swap_in_buffers()}) for {len(params)} parameters"
        for i, p in enumerate(params):
            n = self.id_to_numel[p.ds_id]
            buf = swap_in_buffers[i][:n] if swap_in_buffers is not None else self.get_buffer(p, n)
            assert self.aio_read_handle.async_pread(buf, self._path(p)) == 0
            p.ds_tensor.data = buf
            p.ds_tensor.status = PartitionedParamStatus.INFLIGHT
            self.inflight_params.append(p)
            self.pending_reads += 1
        if not async_op:
            self.synchronize_reads()

    def synchronize_reads(self):
        if self.pending_reads:
            self.aio_read_handle.wait()
            self.pending_reads = 0
        for p in self.inflight_params:
            p.ds_tensor.status = PartitionedParamStatus.AVAILABLE
            self.available_params.add(p.ds_id)
        self.inflight_params = []

    def remove_partition_and_release_buffers(self, params):
        for p in params:
            self._release(p)
