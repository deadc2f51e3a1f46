"""NVTX range decorator (parity: deepspeed/utils/nvtx.py:9-19)."""
import functools

from ..accelerator import get_accelerator


def instrument_w_nvtx(func):

    @functools.wraps(func)
    def wrapped(*args, **kwargs):
        acc = get_accelerator()
        acc.range_push(func.__qualname__)
        try:
            return func(*args, **kwargs)
        finally:
            acc.range_pop()

    return wrapped
