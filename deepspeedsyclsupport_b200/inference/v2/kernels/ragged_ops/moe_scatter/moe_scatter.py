"""(parity: inference/v2/kernels/ragged_ops/moe_scatter/moe_scatter.py)  The kernel objects of this group are defined together in
kernels/ragged_ops/__init__.py; this module keeps the reference import path."""
from .. import MoEScatter  # noqa: F401
