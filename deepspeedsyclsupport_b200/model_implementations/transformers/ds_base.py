"""(parity: deepspeed/model_implementations/transformers/ds_base.py)"""
import torch.nn as nn


class DeepSpeedTransformerBase(nn.Module):
    """Marker base for transformer implementations that are not built from DeepSpeedTransformerInference."""

    def __init__(self):
        super().__init__()
