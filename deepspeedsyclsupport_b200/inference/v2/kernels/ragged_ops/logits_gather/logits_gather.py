"""(parity: inference/v2/kernels/ragged_ops/logits_gather/logits_gather.py)  The kernel objects of this group are defined together in
kernels/ragged_ops/__init__.py; this module keeps the reference import path."""
from .. import RaggedLogitsGather  # noqa: F401
