"""Mixture-of-experts MLP (parity: inference/v2/modules/implementations/moe/cutlass_multi_gemm.py:28-222: top-k gating ->
moe_scatter -> grouped GEMM -> activation -> grouped GEMM -> moe_gather).

The registry key keeps the reference's name; the grouped GEMM is this repo's persistent tcgen05 kernel
(csrc/gemm/gemm_grouped.cu) that reads the per-expert row ranges on the device, so nothing between gating and the combine
synchronises with the host."""
from typing import Any, Dict, Optional

import torch

from ......ops import fused_ops as FO
from ......ops.gemm import moe_gemm
from .... import kernels as K
from ....inference_parameter import InferenceParameter
from ....inference_utils import ActivationType, DtypeEnum, activation_name, is_gated
from ...configs import DSMoEConfig
from ...interfaces import DSMoEBase, DSMoERegistry


@DSMoERegistry.register_module
class DSMultiGemmMoE(DSMoEBase):

    @staticmethod
    def name() -> str:
        return "cutlass_multi_gemm_moe"

    @staticmethod
    def supports_config(config: DSMoEConfig) -> bool:
        if config.input_dtype != config.output_dtype:
            return False
        if DtypeEnum(config.input_dtype) not in (DtypeEnum.fp16, DtypeEnum.bf16, DtypeEnum.fp32):
            return False
        return 1 <= config.top_k <= min(config.n_experts, 4)

    def __init__(self, config: DSMoEConfig, implementation_config: Dict[str, Any] = None) -> None:
        super().__init__(config, implementation_config)
        self._act = ActivationType(config.activation)
        self._gated = is_gated(self._act)
        self._act_name = activation_name(self._act)
        self._output = None

    def transform_gate_param(self, param: torch.Tensor) -> InferenceParameter:
        return InferenceParameter.initialize(param.to(torch.float32).contiguous())  # routing decisions are taken in fp32

    def transform_moe_mlp_1_param(self, param: torch.Tensor) -> InferenceParameter:
        return InferenceParameter.initialize(param.to(DtypeEnum(self._config.input_dtype).value).contiguous())

    def transform_moe_mlp_2_param(self, param: torch.Tensor) -> InferenceParameter:
        return InferenceParameter.initialize(param.to(DtypeEnum(self._config.input_dtype).value).contiguous())

    @property
    def output(self) -> torch.Tensor:
        return self._output

    def forward(self, hidden_states: torch.Tensor, gate_w: torch.Tensor, mlp_1_w: torch.Tensor, mlp_2_w: torch.Tensor,
                mlp_1_b: Optional[torch.Tensor] = None, mlp_2_b: Optional[torch.Tensor] = None) -> torch.Tensor:
        c = self._config
        logits = torch.nn.functional.linear(hidden_states.float(), gate_w)
        counts, scores, assignments, offsets = K.top_k_gating(logits, c.top_k, normalize=c.normalize_scores or c.top_k > 1)
        moe_in, cumsum, slots = K.moe_scatter(hidden_states, counts, assignments, offsets)
        h = moe_gemm(moe_in, mlp_1_w, cumsum)
        if mlp_1_b is not None:
            h = h + self._expand_bias(mlp_1_b, cumsum, h.shape[0])
        h = FO.gated_act(h, self._act_name) if self._gated else FO.bias_act(h, None, self._act_name)
        out = moe_gemm(h, mlp_2_w, cumsum)
        if mlp_2_b is not None:
            out = out + self._expand_bias(mlp_2_b, cumsum, out.shape[0])
        self._output = K.moe_gather(out, scores, slots)
        return self._output

    @staticmethod
    def _expand_bias(bias, cumsum, rows):
        """[E, N] expert biases -> one row per scattered token (rows are sorted by expert)."""
        expert_of_row = torch.searchsorted(cumsum, torch.arange(rows, device=cumsum.device), right=True).clamp_(max=bias.shape[0] - 1)
        return bias[expert_of_row]
