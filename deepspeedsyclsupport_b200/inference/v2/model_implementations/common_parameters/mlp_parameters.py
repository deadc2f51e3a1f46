"""(parity: inference/v2/model_implementations/common_parameters/mlp_parameters.py)"""
import torch

from ..parameter_base import ParameterBase


class MLP1Parameter(ParameterBase):
    """Up projection of an un-gated MLP (weight or bias)."""
    params: torch.Tensor

    def finalize(self) -> torch.Tensor:
        return self.inference_model.transform_mlp_1_param(self.params, gated=False)


class GatedMLPParameter(ParameterBase):
    """gate and up projections, fused as [gate; up] so one GEMM feeds act(gate) * up."""
    gate_params: torch.Tensor
    up_params: torch.Tensor

    def finalize(self) -> torch.Tensor:
        return self.inference_model.transform_mlp_1_param(torch.cat([self.gate_params, self.up_params], dim=0), gated=True)


class MLP2Parameter(ParameterBase):
    """Down projection (weight or bias)."""
    params: torch.Tensor

    def finalize(self) -> torch.Tensor:
        return self.inference_model.transform_mlp_2_param(self.params)
