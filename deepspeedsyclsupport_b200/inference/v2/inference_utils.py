"""Enums and small helpers shared by the FastGen modules (parity: inference/v2/inference_utils.py:12-105)."""
from enum import Enum, IntEnum

import torch


class NormTypeEnum(Enum):
    LayerNorm = "layer_norm"
    RMSNorm = "rms_norm"


class DtypeEnum(Enum):
    """A dtype that can be spelled several ways in a config: DtypeEnum("bf16") is DtypeEnum(torch.bfloat16)."""
    fp16 = torch.float16, "torch.float16", "fp16", "float16", "half"
    fp32 = torch.float32, "torch.float32", "fp32", "float32", "float"
    bf16 = torch.bfloat16, "torch.bfloat16", "bf16", "bfloat16", "bfloat"
    int8 = torch.int8, "torch.int8", "int8"
    fp8 = torch.float8_e4m3fn, "torch.float8_e4m3fn", "fp8", "e4m3"

    def __new__(cls, *values):
        obj = object.__new__(cls)
        obj._value_ = values[0]
        for alias in values[1:]:
            cls._value2member_map_[alias] = obj
        obj._aliases = values
        return obj

    def __repr__(self):
        return f"<{type(self).__name__}.{self._name_}: {', '.join(repr(v) for v in self._aliases)}>"


class ActivationType(IntEnum):
    GELU = 0
    RELU = 1
    SILU = 2
    GEGLU = 3
    ReGLU = 4
    SiGLU = 5
    IDENTITY = 6
    InvalidType = -1


_GATED = (ActivationType.GEGLU, ActivationType.ReGLU, ActivationType.SiGLU)
# name understood by ops/fused_ops (bias_act / gated_act)
_ACT_NAME = {ActivationType.GELU: "gelu", ActivationType.RELU: "relu", ActivationType.SILU: "silu", ActivationType.GEGLU: "gelu",
             ActivationType.ReGLU: "relu", ActivationType.SiGLU: "silu", ActivationType.IDENTITY: "identity"}


def is_gated(act_fn) -> bool:
    return ActivationType(act_fn) in _GATED


def activation_name(act_fn) -> str:
    return _ACT_NAME[ActivationType(act_fn)]


def elem_size(dtype: torch.dtype) -> int:
    if not isinstance(dtype, torch.dtype):
        raise ValueError(f"Unknown dtype size for {dtype}")
    return torch.empty((), dtype=dtype).element_size()


def ceil_div(a: int, b: int) -> int:
    return -(-a // b)
