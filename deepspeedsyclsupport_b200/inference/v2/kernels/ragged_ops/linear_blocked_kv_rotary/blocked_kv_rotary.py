"""(parity: inference/v2/kernels/ragged_ops/linear_blocked_kv_rotary/blocked_kv_rotary.py)  The kernel objects of this group are defined together in
kernels/ragged_ops/__init__.py; this module keeps the reference import path."""
from .. import BlockedRotaryEmbeddings  # noqa: F401
