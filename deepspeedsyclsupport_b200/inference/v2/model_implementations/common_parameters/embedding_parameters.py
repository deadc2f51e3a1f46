"""(parity: inference/v2/model_implementations/common_parameters/embedding_parameters.py)"""
import torch

from ..parameter_base import ParameterBase


class EmbeddingParameter(ParameterBase):
    """Token or learned-position embedding table."""
    params: torch.Tensor

    def finalize(self) -> torch.Tensor:
        return self.inference_model.transform_embedding_param(self.params)
