"""(parity: inference/v2/kernels/cutlass_ops/moe_gemm/moe_gemm.py)  The kernel objects of this group are defined together in
kernels/cutlass_ops/__init__.py; this module keeps the reference import path."""
from .. import MoEGEMM  # noqa: F401
