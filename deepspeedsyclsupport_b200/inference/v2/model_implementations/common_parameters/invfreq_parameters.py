"""(parity: inference/v2/model_implementations/common_parameters/invfreq_parameters.py)"""
import torch

from ..parameter_base import ParameterBase


class InvFreqParameter(ParameterBase):
    """Trained rotary inverse frequencies stored in the checkpoint (kept in fp32, replicated)."""
    params: torch.Tensor

    def finalize(self) -> torch.Tensor:
        return self.inference_model.transform_inv_freq_param(self.params)
