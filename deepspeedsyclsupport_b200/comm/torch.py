"""`TorchBackend`: object view of the communication layer (reference: deepspeed/comm/torch.py:99-368).  Every method forwards
to the module-level op of `comm.comm` (which adds timing / logging and the shared-memory and gloo special cases), so both
spellings behave identically."""
import torch.distributed as tdist

from . import comm as _c
from .backend import Backend


def has_coalescing_manager():
    return hasattr(tdist.distributed_c10d, "_coalescing_manager")


def has_all_reduce_coalesced():
    return hasattr(tdist, "all_reduce_coalesced")


class TorchBackend(Backend):

    def __init__(self, backend, timeout=None, init_method=None, rank=-1, world_size=-1, name="torch"):
        super().__init__(name=name)
        self.has_all_reduce_coalesced = has_all_reduce_coalesced()
        self.has_coalescing_manager = has_coalescing_manager()
        self.all_gather_function = self.get_all_gather_function()
        self.reduce_scatter_function = self.get_reduce_scatter_function()
        self.single_gpu_mode = True
        self.backend_name = backend
        if not tdist.is_initialized():
            self.init_process_group(backend, timeout, init_method, rank, world_size)
        self.initialized = True
        self.using_mpi = False

    @classmethod
    def get_all_gather_function(cls):
        return getattr(tdist, "all_gather_into_tensor", getattr(tdist, "_all_gather_base", None))

    @classmethod
    def get_reduce_scatter_function(cls):
        return getattr(tdist, "reduce_scatter_tensor", getattr(tdist, "_reduce_scatter_base", None))

    def has_all_gather_into_tensor(self):
        return self.all_gather_function is not None

    def has_reduce_scatter_tensor(self):
        return self.reduce_scatter_function is not None

    def init_process_group(self, backend, timeout=None, init_method=None, rank=-1, world_size=-1):
        if not tdist.is_initialized():
            _c.init_distributed(dist_backend=backend, init_method=init_method, rank=rank, world_size=world_size,
                                **({"timeout": timeout} if timeout is not None else {}))
        self.using_mpi = tdist.get_backend() == "mpi"

    def all_gather_base(self, output_tensor, input_tensor, group=None, async_op=False):
        return _c.all_gather_into_tensor(output_tensor, input_tensor, group=group, async_op=async_op)

    def _reduce_op(self, op):
        return _c._op(op)


def _forward(name):

    def method(self, *args, **kwargs):
        return getattr(_c, name)(*args, **kwargs)

    method.__name__ = name
    return method


for _name in ("all_reduce", "inference_all_reduce", "all_reduce_coalesced", "reduce", "reduce_scatter", "broadcast", "all_gather",
              "all_gather_into_tensor", "all_gather_coalesced", "reduce_scatter_tensor", "all_to_all_single", "all_to_all", "send", "recv",
              "isend", "irecv", "gather", "scatter", "barrier", "monitored_barrier", "get_rank", "get_world_size", "is_initialized",
              "get_backend", "new_group", "get_global_rank", "get_world_group", "destroy_process_group"):
    if hasattr(_c, _name):
        setattr(TorchBackend, _name, _forward(_name))
