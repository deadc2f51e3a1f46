"""(parity: inference/v2/model_implementations/common_parameters/norm_parameters.py)"""
import torch

from ..parameter_base import ParameterBase


class NormParameter(ParameterBase):
    """gamma or beta of a LayerNorm / RMSNorm (replicated on every tensor-parallel rank)."""
    params: torch.Tensor

    def finalize(self) -> torch.Tensor:
        return self.inference_model.transform_norm_param(self.params)
