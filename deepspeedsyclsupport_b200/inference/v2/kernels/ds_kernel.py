"""(parity: inference/v2/kernels/ds_kernel.py:9-32)"""
from abc import ABC, abstractmethod


class DSKernelBase(ABC):
    """A kernel object is constructed once per configuration (dtype, widths, activation): the constructor validates what the
    device code supports and raises ValueError otherwise; `__call__` only launches.  The FastGen modules hold kernel objects so a
    bad configuration fails at model build time, not inside a forward."""

    @abstractmethod
    def __init__(self, *args, **kwargs):
        raise NotImplementedError()

    @abstractmethod
    def __call__(self, *args, **kwargs):
        raise NotImplementedError()
