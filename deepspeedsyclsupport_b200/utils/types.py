"""Enumerations shared by the inference kernels and model implementations: which activation a fused MLP applies and
which normalisation a block uses.  Values are wire-stable (they are passed to native kernels as ints)."""
import enum

ActivationFuncType = enum.IntEnum("ActivationFuncType", {"UNKNOWN": 0, "GELU": 1, "ReLU": 2, "GATED_GELU": 3, "GATED_SILU": 4})
NormType = enum.IntEnum("NormType", {"UNKNOWN": 0, "LayerNorm": 1, "GroupNorm": 2, "RMSNorm": 3})

# activations whose up-projection produces [gate | value] pairs
GATED_ACTIVATION_TYPES = [ActivationFuncType.GATED_GELU, ActivationFuncType.GATED_SILU]
