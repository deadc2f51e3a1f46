"""(parity: inference/v2/model_implementations/phi/containers.py -- the reference names this family's module in the plural)"""
from .container import PhiNonTransformerContainer, PhiTransformerContainer  # noqa: F401
