"""Accelerator interface kept as an API shim.

The reference selects among cuda/cpu/xpu/npu/mps/hpu at run time
(accelerator/abstract_accelerator.py:10-277, real_accelerator.py:51-192).  This build has exactly
one device backend -- CUDA on sm_100a -- so the ABC survives only because user code and the
reference's tests call `get_accelerator()`; a host ("cpu") flavour of the same class exists so the
plumbing tests run with gloo on machines without a GPU.  There is no op dispatch through it.
"""
import abc


class DeepSpeedAccelerator(abc.ABC):

    def __init__(self):
        self._name = None
        self._communication_backend_name = None

    @abc.abstractmethod
    def is_synchronized_device(self):
        ...

    @abc.abstractmethod
    def device_name(self, device_index=None):
        ...

    @abc.abstractmethod
    def device(self, device_index=None):
        ...

    @abc.abstractmethod
    def set_device(self, device_index):
        ...

    @abc.abstractmethod
    def current_device(self):
        ...

    @abc.abstractmethod
    def current_device_name(self):
        ...

    @abc.abstractmethod
    def device_count(self):
        ...

    @abc.abstractmethod
    def synchronize(self, device_index=None):
        ...

    @abc.abstractmethod
    def communication_backend_name(self):
        ...
