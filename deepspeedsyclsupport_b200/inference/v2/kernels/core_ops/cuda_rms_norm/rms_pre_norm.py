"""(parity: inference/v2/kernels/core_ops/cuda_rms_norm/rms_pre_norm.py)  The kernel objects of this group are defined together in
kernels/core_ops/__init__.py; this module keeps the reference import path."""
from .. import CUDARMSPreNorm  # noqa: F401
