"""(parity: inference/v2/kernels/ragged_ops/embed/embed.py)  The kernel objects of this group are defined together in
kernels/ragged_ops/__init__.py; this module keeps the reference import path."""
from .. import RaggedEmbeddingKernel  # noqa: F401
