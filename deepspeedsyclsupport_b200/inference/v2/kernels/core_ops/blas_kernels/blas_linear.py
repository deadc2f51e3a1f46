"""(parity: inference/v2/kernels/core_ops/blas_kernels/blas_linear.py)  The kernel objects of this group are defined together in
kernels/core_ops/__init__.py; this module keeps the reference import path."""
from .. import BlasLibLinear  # noqa: F401
