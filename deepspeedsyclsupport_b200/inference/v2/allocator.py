"""Device placement helpers for model-building code (parity: inference/v2/allocator.py:13-32).

`on_device` wraps a function that returns a tensor (or nothing) and moves the result to the current accelerator device, keeping
the InferenceParameter metadata; `empty_from` carves a shaped view out of a pre-allocated buffer (the FastGen modules reuse
one output buffer across forwards)."""
import functools
from functools import reduce
from typing import Iterable

import torch

from ...accelerator import get_accelerator


def on_device(method):

    @functools.wraps(method)
    def wrapped(self, *args, **kwargs):
        ret = method(self, *args, **kwargs)
        if isinstance(ret, torch.Tensor):
            from .inference_parameter import InferenceParameter
            moved = ret.to(get_accelerator().current_device_name())
            if isinstance(ret, InferenceParameter):
                return InferenceParameter.initialize(moved, **ret.aux_attrs)
            return moved
        return ret

    return wrapped


def empty_from(tensor: torch.Tensor, shape: Iterable[int]) -> torch.Tensor:
    shape = tuple(shape)
    n = reduce(lambda x, y: x * y, shape, 1)
    assert n <= tensor.numel(), f"buffer of {tensor.numel()} elements cannot hold shape {shape}"
    return tensor.flatten()[:n].view(shape)
