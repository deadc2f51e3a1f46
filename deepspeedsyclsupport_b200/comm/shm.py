"""Single-node all-reduce between host processes through POSIX shared memory (csrc/host/shm_comm.cpp).

`inference_all_reduce` on host tensors goes through this instead of gloo when every rank of the group runs on one host:
each rank copies its chunk into its slot of the segment, reduces 1/world of the elements over all slots and copies the
result out -- two process-shared barriers per 1 MiB chunk, no sockets.
(reference: csrc/cpu/comm/ccl.cpp:540-611 `inference_all_reduce`, comm/ccl.py:61-76)"""
import ctypes
import os
import socket
import uuid

import torch
import torch.distributed as tdist

_DTYPES = {torch.float32: 0, torch.bfloat16: 1, torch.float16: 2, torch.int32: 3}
_handles = {}  # group key -> ShmAllReduce or None (probed, not usable)
CHUNK_BYTES = 1 << 20
TIMEOUT_S = float(os.environ.get("DSB200_SHM_TIMEOUT", "300"))


def _host_lib():
    from ..ops import native
    l = native.host_lib()
    if not getattr(l, "_shm_sigs", False):
        l.dsbh_shm_open.restype = ctypes.c_void_p
        l.dsbh_shm_open.argtypes = [ctypes.c_char_p, ctypes.c_int, ctypes.c_int, ctypes.c_longlong, ctypes.c_double]
        l.dsbh_shm_allreduce.restype = ctypes.c_int
        l.dsbh_shm_allreduce.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_longlong, ctypes.c_int, ctypes.c_double]
        l.dsbh_shm_barrier.restype = ctypes.c_int
        l.dsbh_shm_barrier.argtypes = [ctypes.c_void_p, ctypes.c_double]
        l.dsbh_shm_close.restype = None
        l.dsbh_shm_close.argtypes = [ctypes.c_void_p]
        l._shm_sigs = True
    return l


class ShmAllReduce(object):
    """One shared segment for one process group.  Construction is collective over the group."""

    def __init__(self, group=None, chunk_bytes=CHUNK_BYTES):
        self.group = group
        self.rank = tdist.get_rank(group)
        self.world = tdist.get_world_size(group)
        token = [f"dsb200_{os.getpid()}_{uuid.uuid4().hex[:12]}" if self.rank == 0 else None]
        tdist.broadcast_object_list(token, src=tdist.get_global_rank(group, 0) if group is not None else 0, group=group)
        self._lib = _host_lib()
        self._h = self._lib.dsbh_shm_open(token[0].encode(), self.rank, self.world, chunk_bytes, TIMEOUT_S)
        if not self._h:
            raise RuntimeError(f"shared-memory segment {token[0]} could not be opened on rank {self.rank}")

    def all_reduce(self, tensor):
        assert tensor.device.type == "cpu" and tensor.is_contiguous() and tensor.dtype in _DTYPES
        rc = self._lib.dsbh_shm_allreduce(self._h, tensor.data_ptr(), tensor.numel(), _DTYPES[tensor.dtype], TIMEOUT_S)
        if rc == 2:
            raise RuntimeError("shared-memory all-reduce: a peer did not arrive (dead rank?)")
        if rc != 0:
            raise RuntimeError("shared-memory all-reduce: bad argument")
        return tensor

    def barrier(self):
        if self._lib.dsbh_shm_barrier(self._h, TIMEOUT_S) != 0:
            raise RuntimeError("shared-memory barrier: a peer did not arrive")

    def close(self):
        if self._h:
            self._lib.dsbh_shm_close(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def _single_host(group):
    names = [None] * tdist.get_world_size(group)
    tdist.all_gather_object(names, socket.gethostname(), group=group)
    return len(set(names)) == 1


def for_group(group=None):
    """The group's segment, created on first use (collective); None when the group spans hosts or has one rank."""
    key = id(group) if group is not None else 0
    if key not in _handles:
        h = None
        if tdist.get_world_size(group) > 1 and os.environ.get("DSB200_DISABLE_SHM", "0") != "1" and _single_host(group):
            h = ShmAllReduce(group)
        _handles[key] = h
    return _handles[key]


def usable(tensor, op_is_sum):
    return (op_is_sum and tensor.device.type == "cpu" and tensor.is_contiguous() and tensor.dtype in _DTYPES)


def shutdown():
    for h in _handles.values():
        if h is not None:
            h.close()
    _handles.clear()
