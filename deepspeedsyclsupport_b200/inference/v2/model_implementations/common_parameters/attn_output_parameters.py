"""(parity: inference/v2/model_implementations/common_parameters/attn_output_parameters.py)"""
from .attn_parameters import AttentionOutputParameter  # noqa: F401
