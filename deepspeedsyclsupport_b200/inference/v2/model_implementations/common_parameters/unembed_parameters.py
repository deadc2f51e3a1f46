"""(parity: inference/v2/model_implementations/common_parameters/unembed_parameters.py)"""
import torch

from ..parameter_base import ParameterBase


class UnembedParameter(ParameterBase):
    """Output projection onto the vocabulary (weight [vocab, hidden] or bias [vocab]); vocabulary-parallel."""
    params: torch.Tensor

    def finalize(self) -> torch.Tensor:
        return self.inference_model.transform_unembed_param(self.params)
