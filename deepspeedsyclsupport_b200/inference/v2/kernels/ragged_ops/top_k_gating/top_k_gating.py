"""(parity: inference/v2/kernels/ragged_ops/top_k_gating/top_k_gating.py)  The kernel objects of this group are defined together in
kernels/ragged_ops/__init__.py; this module keeps the reference import path."""
from .. import RaggedTopKGating  # noqa: F401
