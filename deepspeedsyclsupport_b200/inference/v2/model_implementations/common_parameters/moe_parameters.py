"""(parity: inference/v2/model_implementations/common_parameters/moe_parameters.py)  Expert matrices arrive one per expert
and leave stacked [E, out, in]: the grouped GEMM takes one base pointer and finds expert e at a fixed stride."""
import torch

from ..parameter_base import ParameterBase, ParamList


class MoEGatingWeightParameter(ParameterBase):
    """Router [n_experts, hidden] (replicated; fp32)."""
    params: torch.Tensor

    def finalize(self) -> torch.Tensor:
        return self.inference_model.transform_moe_gate_param(self.params)


class UnfusedMoEMLP1Parameter(ParameterBase):
    experts: ParamList("n_experts")

    def finalize(self) -> torch.Tensor:
        return self.inference_model.transform_moe_mlp_1_param(torch.stack(list(self.experts), dim=0), gated=False)


class UnfusedMoEGatedMLPParameter(ParameterBase):
    gating_experts: ParamList("n_experts")
    up_experts: ParamList("n_experts")

    def finalize(self) -> torch.Tensor:
        fused = torch.stack([torch.cat([g, u], dim=0) for g, u in zip(self.gating_experts, self.up_experts)], dim=0)
        return self.inference_model.transform_moe_mlp_1_param(fused, gated=True)


class UnfusedMoEMLP2Parameter(ParameterBase):
    experts: ParamList("n_experts")

    def finalize(self) -> torch.Tensor:
        return self.inference_model.transform_moe_mlp_2_param(torch.stack(list(self.experts), dim=0))
