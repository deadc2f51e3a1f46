"""(parity: inference/v2/kernels/ragged_ops/moe_gather/moe_gather.py)  The kernel objects of this group are defined together in
kernels/ragged_ops/__init__.py; this module keeps the reference import path."""
from .. import MoEGather  # noqa: F401
