"""Host-side backend for CPU tensor-parallel inference (parity: deepspeed/comm/ccl.py:15-193, whose oneCCL build is an
Intel-CPU component).  On this platform the same role -- low-latency collectives between the host processes of one node -- is
played by gloo for the general case plus the POSIX shared-memory all-reduce of csrc/host/shm_comm.cpp for the hot path
(`inference_all_reduce`).  `CCLBackend` is a TorchBackend over gloo whose all_reduce / inference_all_reduce try the
shared-memory segment first."""
import torch

from . import shm
from .reduce_op import ReduceOp
from .torch import TorchBackend


def build_ccl_op():
    """The native library that implements the shared-memory path, or None if it is not built."""
    try:
        from ..ops import native
        return native.host_lib()
    except Exception:
        return None


class CCLHandler():
    """Completed-work handle (the shared-memory all-reduce is synchronous)."""

    def __init__(self, ccl_comm_op=None):
        self.ccl_comm_op = ccl_comm_op

    def wait(self):
        pass

    def is_completed(self):
        return True


class CCLBackend(TorchBackend):

    def __init__(self, name="ccl", rank=-1, world_size=-1, mpu=None, timeout=None, init_method=None):
        self.ccl_comm_op = build_ccl_op()
        if self.ccl_comm_op is None:
            self.initialized = False
            return
        super().__init__(backend="gloo", name="torch", rank=rank, world_size=world_size, timeout=timeout, init_method=init_method)
        self.name = "ccl"
        self.available_coll = self.get_all_available_coll()

    def is_initialized(self):
        return bool(getattr(self, "initialized", False))

    def get_all_available_coll(self):
        return ["broadcast", "all_reduce", "inference_all_reduce", "all_reduce_caching", "barrier"]

    def run_collective(self, name, **kwargs):
        if name in self.available_coll:
            return getattr(self, name)(**kwargs) or CCLHandler(self.ccl_comm_op)
        return getattr(super(), name)(**kwargs)

    def _shm_all_reduce(self, tensor, op, group):
        if op in (ReduceOp.SUM, torch.distributed.ReduceOp.SUM) and shm.usable(tensor, True):
            seg = shm.for_group(group)
            if seg is not None:
                seg.all_reduce(tensor)
                return True
        return False

    def all_reduce(self, tensor, op=ReduceOp.SUM, group=None, async_op=False):
        if not async_op and self._shm_all_reduce(tensor, op, group):
            return CCLHandler(self.ccl_comm_op)
        return super().all_reduce(tensor, op=op, group=group, async_op=async_op)

    def inference_all_reduce(self, tensor, op=ReduceOp.SUM, group=None, async_op=False):
        return self.all_reduce(tensor, op=op, group=group, async_op=async_op)

    def all_reduce_caching(self, tensor, op=ReduceOp.SUM, match_id=None, group=None, async_op=False):
        """The reference caches the communicator lookup per `match_id`; the shared-memory segment is already cached per group."""
        return self.all_reduce(tensor, op=op, group=group, async_op=async_op)
