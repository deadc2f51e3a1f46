"""(parity: inference/v2/kernels/core_ops/cuda_layer_norm/cuda_ln.py)  The kernel objects of this group are defined together in
kernels/core_ops/__init__.py; this module keeps the reference import path."""
from .. import CUDAFPLN  # noqa: F401
