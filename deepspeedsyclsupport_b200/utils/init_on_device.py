"""OnDevice: build a model directly on a target device / floating dtype (typically `meta`, to describe a model without
allocating it, or `cuda` + bf16 to skip the fp32 CPU copy).

Implemented with a TorchFunctionMode: every factory call made while the context is active (torch.empty / zeros / ones /
full / rand / randn / arange / tensor and the nn.init-time `Parameter(torch.empty(...))` pattern) gets `device=` filled in
when the caller did not pass one, and floating-point results are cast to the requested dtype.  This covers the
reference's patched constructors (deepspeed/utils/init_on_device.py:12-81) without monkey-patching torch."""
import torch
from torch.overrides import TorchFunctionMode

_FACTORIES = {
    torch.empty, torch.zeros, torch.ones, torch.full, torch.rand, torch.randn, torch.arange, torch.tensor, torch.eye, torch.linspace,
    torch.empty_strided, torch.randint, torch.randperm, torch.as_tensor,
}


class _FactoryMode(TorchFunctionMode):

    def __init__(self, device, dtype):
        super().__init__()
        self.device, self.dtype = device, dtype

    def __torch_function__(self, func, types, args=(), kwargs=None):
        kwargs = kwargs or {}
        if func in _FACTORIES:
            if kwargs.get("device") is None:
                kwargs["device"] = self.device
            out = func(*args, **kwargs)
            if self.dtype is not None and torch.is_tensor(out) and out.is_floating_point() and out.dtype != self.dtype:
                out = out.to(self.dtype)
            return out
        return func(*args, **kwargs)


class OnDevice(object):
    """with OnDevice(dtype=torch.bfloat16, device="meta"): model = Model()"""

    def __init__(self, dtype, device="meta", enabled=True):
        self.dtype = dtype
        self.device = device
        self.enabled = enabled
        self._mode = None

    def __enter__(self):
        if self.enabled:
            self._mode = _FactoryMode(torch.device(self.device), self.dtype)
            self._mode.__enter__()
        return self

    def __exit__(self, exc_type, exc_value, traceback):
        if self._mode is not None:
            self._mode.__exit__(exc_type, exc_value, traceback)
            self._mode = None
        return False
