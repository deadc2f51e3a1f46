"""`get_accelerator()` -- the single-backend implementation (see abstract_accelerator.py).

Method surface follows accelerator/abstract_accelerator.py:10-277 and cuda_accelerator.py:22-338 of the
reference: device/RNG/stream/event/memory/dtype/graph helpers, tensor constructors, pin_memory, op-builder
reflection (`create_op_builder/get_op_builder`), `communication_backend_name`, `export_envs`.
"""
import contextlib
import os

import torch

from .abstract_accelerator import DeepSpeedAccelerator

SUPPORTED_ACCELERATOR_LIST = ["cuda", "cpu"]


class B200Accelerator(DeepSpeedAccelerator):
    """CUDA (sm_100a) accelerator; degrades to host tensors + gloo when no GPU is visible."""

    def __init__(self, force_cpu=False):
        super().__init__()
        self._cuda = torch.cuda.is_available() and not force_cpu
        self._name = "cuda" if self._cuda else "cpu"
        self._communication_backend_name = "nccl" if self._cuda else "gloo"
        self._graphs = {}

    # ---- device
    def is_synchronized_device(self):
        return not self._cuda

    def use_host_timers(self):
        return self.is_synchronized_device()

    def resolves_data_dependency(self):
        return self.is_synchronized_device()

    def handles_memory_backpressure(self):
        return self.is_synchronized_device()

    def device_name(self, device_index=None):
        if not self._cuda:
            return "cpu"
        return "cuda" if device_index is None else f"cuda:{device_index}"

    def device(self, device_index=None):
        return torch.cuda.device(device_index) if self._cuda else contextlib.nullcontext()

    def set_device(self, device_index):
        if self._cuda:
            torch.cuda.set_device(device_index)

    def current_device(self):
        return torch.cuda.current_device() if self._cuda else 0

    def current_device_name(self):
        return f"cuda:{torch.cuda.current_device()}" if self._cuda else "cpu"

    def device_count(self):
        return torch.cuda.device_count() if self._cuda else int(os.environ.get("LOCAL_SIZE", "1"))

    def synchronize(self, device_index=None):
        if self._cuda:
            torch.cuda.synchronize(device_index)

    # ---- RNG
    def random(self):
        return torch.random

    def set_rng_state(self, new_state, device_index=None):
        if self._cuda:
            return torch.cuda.set_rng_state(new_state) if device_index is None else torch.cuda.set_rng_state(new_state, device_index)
        return torch.set_rng_state(new_state)

    def get_rng_state(self, device_index=None):
        if self._cuda:
            return torch.cuda.get_rng_state() if device_index is None else torch.cuda.get_rng_state(device_index)
        return torch.get_rng_state()

    def manual_seed(self, seed):
        return torch.cuda.manual_seed(seed) if self._cuda else torch.manual_seed(seed)

    def manual_seed_all(self, seed):
        return torch.cuda.manual_seed_all(seed) if self._cuda else torch.manual_seed(seed)

    def initial_seed(self, seed=None):
        return torch.cuda.initial_seed() if self._cuda else torch.initial_seed()

    def default_generator(self, device_index):
        return torch.cuda.default_generators[device_index] if self._cuda else torch.default_generator

    # ---- streams / events
    @property
    def Stream(self):
        return torch.cuda.Stream if self._cuda else None

    def stream(self, stream):
        return torch.cuda.stream(stream) if self._cuda and stream is not None else contextlib.nullcontext()

    def current_stream(self, device_index=None):
        return torch.cuda.current_stream(device_index) if self._cuda else None

    def default_stream(self, device_index=None):
        return torch.cuda.default_stream(device_index) if self._cuda else None

    @property
    def Event(self):
        return torch.cuda.Event if self._cuda else None

    # ---- memory
    def empty_cache(self):
        if self._cuda:
            torch.cuda.empty_cache()

    def _mem(self, fn, device_index=None):
        if self._cuda:
            return getattr(torch.cuda, fn)(device_index)
        try:
            import psutil
            return psutil.Process().memory_info().rss
        except Exception:
            return 0

    def memory_allocated(self, device_index=None):
        return self._mem("memory_allocated", device_index)

    def max_memory_allocated(self, device_index=None):
        return self._mem("max_memory_allocated", device_index)

    def reset_max_memory_allocated(self, device_index=None):
        if self._cuda:
            torch.cuda.reset_peak_memory_stats(device_index)

    def memory_cached(self, device_index=None):
        return self._mem("memory_reserved", device_index)

    def max_memory_cached(self, device_index=None):
        return self._mem("max_memory_reserved", device_index)

    def reset_max_memory_cached(self, device_index=None):
        if self._cuda:
            torch.cuda.reset_peak_memory_stats(device_index)

    def memory_stats(self, device_index=None):
        return torch.cuda.memory_stats(device_index) if self._cuda else {}

    def reset_peak_memory_stats(self, device_index=None):
        if self._cuda:
            torch.cuda.reset_peak_memory_stats(device_index)

    def memory_reserved(self, device_index=None):
        return self._mem("memory_reserved", device_index)

    def max_memory_reserved(self, device_index=None):
        return self._mem("max_memory_reserved", device_index)

    def total_memory(self, device_index=None):
        if self._cuda:
            return torch.cuda.get_device_properties(device_index or 0).total_memory
        import psutil
        return psutil.virtual_memory().total

    def available_memory(self, device_index=None):
        if self._cuda:
            free, _ = torch.cuda.mem_get_info(device_index)
            return free
        import psutil
        return psutil.virtual_memory().available

    # ---- dtype support (B200: everything incl. fp8)
    def is_bf16_supported(self):
        return True

    def is_fp16_supported(self):
        return True

    def is_fp8_supported(self):
        return self._cuda

    def supported_dtypes(self):
        return [torch.float, torch.half, torch.bfloat16]

    # ---- misc
    def amp(self):
        return torch.cuda.amp if self._cuda else None

    def is_available(self):
        return True

    def range_push(self, msg):
        if self._cuda:
            return torch.cuda.nvtx.range_push(msg)

    def range_pop(self):
        if self._cuda:
            return torch.cuda.nvtx.range_pop()

    def lazy_call(self, callback):
        return torch.cuda._lazy_call(callback) if self._cuda else callback()

    def communication_backend_name(self):
        return self._communication_backend_name

    def is_triton_supported(self):
        return False  # by design: no Triton anywhere in this framework

    # ---- CUDA graphs
    def create_graph(self):
        return torch.cuda.CUDAGraph() if self._cuda else None

    def capture_to_graph(self, graph, pool=None, stream=None):
        return torch.cuda.graph(graph, pool, stream) if self._cuda else contextlib.nullcontext()

    def replay_graph(self, graph):
        if graph is not None:
            graph.replay()

    # ---- tensor constructors
    def _tt(self, dtype):
        import functools
        return functools.partial(torch.tensor, dtype=dtype, device=self.current_device_name())

    @property
    def BFloat16Tensor(self):
        return self._tt(torch.bfloat16)

    @property
    def ByteTensor(self):
        return self._tt(torch.uint8)

    @property
    def DoubleTensor(self):
        return self._tt(torch.double)

    @property
    def FloatTensor(self):
        return self._tt(torch.float)

    @property
    def HalfTensor(self):
        return self._tt(torch.half)

    @property
    def IntTensor(self):
        return self._tt(torch.int)

    @property
    def LongTensor(self):
        return self._tt(torch.long)

    def pin_memory(self, tensor, align_bytes=1):
        return tensor.pin_memory() if self._cuda else tensor

    def is_pinned(self, tensor):
        return tensor.is_pinned()

    def on_accelerator(self, tensor):
        return tensor.is_cuda if self._cuda else True

    # ---- op builders (reflection API used by user code: get_accelerator().create_op_builder("FusedAdamBuilder"))
    def op_builder_dir(self):
        return "deepspeedsyclsupport_b200.ops.op_builder"

    def create_op_builder(self, class_name):
        cls = self.get_op_builder(class_name)
        return cls() if cls is not None else None

    def get_op_builder(self, class_name):
        from ..ops.op_builder import all_ops
        return all_ops.BUILDERS.get(class_name)

    def build_extension(self):
        from torch.utils.cpp_extension import BuildExtension
        return BuildExtension

    def export_envs(self):
        return ["NCCL", "CUDA", "DSB200"]


ds_accelerator = None


def is_current_accelerator_supported():
    return get_accelerator().device_name() in SUPPORTED_ACCELERATOR_LIST


def get_accelerator():
    global ds_accelerator
    if ds_accelerator is None:
        force_cpu = os.environ.get("DS_ACCELERATOR", "") == "cpu"
        ds_accelerator = B200Accelerator(force_cpu=force_cpu)
    return ds_accelerator


def set_accelerator(accel_obj):
    global ds_accelerator
    assert isinstance(accel_obj, DeepSpeedAccelerator)
    ds_accelerator = accel_obj
