"""(parity: inference/v2/kernels/core_ops/bias_activations/bias_activation.py)  The kernel objects of this group are defined together in
kernels/core_ops/__init__.py; this module keeps the reference import path."""
from .. import CUDABiasActivation  # noqa: F401
