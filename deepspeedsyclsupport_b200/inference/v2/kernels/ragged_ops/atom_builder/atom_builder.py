"""(parity: inference/v2/kernels/ragged_ops/atom_builder/atom_builder.py)  The kernel objects of this group are defined together in
kernels/ragged_ops/__init__.py; this module keeps the reference import path."""
from .. import AtomBuilder  # noqa: F401
