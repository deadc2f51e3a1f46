"""Tensor subclass that carries auxiliary tensors (scales of a quantised weight, ...) through `.to()` and serialisation
(parity: inference/v2/inference_parameter.py:14-89)."""
from typing import Dict

import torch

CORE_PARAM = "_ds_core_param_key"
STR_TO_DTYPE = {str(d): d for d in (torch.float16, torch.bfloat16, torch.float32, torch.int8, torch.uint8, torch.int32, torch.int64,
                                    torch.float8_e4m3fn, torch.bool)}


class InferenceParameter(torch.Tensor):
    # like nn.Parameter: operations on a parameter return plain tensors and skip the Python-level subclass dispatch, so holding
    # weights in this class costs nothing on the forward path
    __torch_function__ = torch._C._disabled_torch_function_impl

    @staticmethod
    def __new__(cls, tensor, *args, **kwargs):
        new_tensor = torch.Tensor._make_subclass(cls, tensor.detach() if tensor.requires_grad else tensor, False)
        if hasattr(tensor, "_aux_attrs"):
            setattr(new_tensor, "_aux_attrs", tensor.aux_attrs)
        return new_tensor

    def to(self, *args, **kwargs):
        """Like Tensor.to for the core tensor; auxiliary tensors follow a device move but keep their own dtype."""
        device, _, non_blocking, _ = torch._C._nn._parse_to(*args, **kwargs)
        core = self.as_subclass(torch.Tensor).to(*args, **kwargs)
        aux = {k: (v.to(device, non_blocking=non_blocking) if device is not None else v) for k, v in self.aux_attrs.items()}
        return InferenceParameter.initialize(core, **aux)

    @classmethod
    def initialize(cls, core_param: torch.Tensor, **kwargs) -> "InferenceParameter":
        param = InferenceParameter(core_param)
        setattr(param, "_aux_attrs", kwargs)
        for attr_name, attr in kwargs.items():
            if hasattr(param, attr_name):
                raise ValueError(f"Attribute {attr_name} already exists on param.")
            if not isinstance(attr, torch.Tensor):
                raise ValueError(f"Attribute {attr_name} must be a tensor.")
            setattr(param, attr_name, attr)
        return param

    @classmethod
    def initialize_raw(cls, **kwargs) -> "InferenceParameter":
        """From a flat dict holding the core tensor under CORE_PARAM (the form used by the flattened-model files)."""
        if CORE_PARAM not in kwargs:
            raise ValueError(f"Must provide core parameter, with key {CORE_PARAM}.")
        return InferenceParameter.initialize(kwargs[CORE_PARAM], **{k: v for k, v in kwargs.items() if k != CORE_PARAM})

    @property
    def aux_attrs(self) -> Dict[str, torch.Tensor]:
        return getattr(self, "_aux_attrs", {})
