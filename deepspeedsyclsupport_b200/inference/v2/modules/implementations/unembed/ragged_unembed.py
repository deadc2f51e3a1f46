"""(parity: inference/v2/modules/implementations/unembed/ragged_unembed.py:24-120: logits_gather + final norm + projection)"""
from typing import Any, Dict, Optional

import torch

from ......ops import fused_ops as FO
from ......ops.gemm import gemm
from .... import kernels as K
from ....inference_utils import DtypeEnum, NormTypeEnum
from ...configs import DSUnembedConfig
from ...interfaces import DSUnembedBase, DSUnembedRegistry


@DSUnembedRegistry.register_module
class DSRaggedUnembed(DSUnembedBase):

    @staticmethod
    def name() -> str:
        return "ragged_unembed"

    @staticmethod
    def supports_config(config: DSUnembedConfig) -> bool:
        return DtypeEnum(config.dtype) in (DtypeEnum.fp16, DtypeEnum.bf16, DtypeEnum.fp32)

    def __init__(self, config: DSUnembedConfig, implementation_config: Dict[str, Any] = None) -> None:
        super().__init__(config, implementation_config)
        self._norm = None if config.norm_type is None else NormTypeEnum(config.norm_type)
        self._output = None

    @property
    def output(self) -> torch.Tensor:
        return self._output

    def forward(self, hidden_states: torch.Tensor, vocab_embedding: torch.Tensor, ragged_metadata, bias: Optional[torch.Tensor] = None,
                gamma: Optional[torch.Tensor] = None, beta: Optional[torch.Tensor] = None,
                residual: Optional[torch.Tensor] = None) -> torch.Tensor:
        """`residual` (optional): the residual stream the last hidden state still has to be added to -- only the gathered rows
        (one per sequence) are added and normalised, not the whole batch."""
        last_tok = ragged_metadata.last_tokens()
        h = K.logits_gather(hidden_states, last_tok)
        r = K.logits_gather(residual, last_tok) if residual is not None else None
        if self._norm == NormTypeEnum.RMSNorm:
            h = FO.rms_norm(h, gamma, self._config.eps, r)
        elif self._norm == NormTypeEnum.LayerNorm:
            h = FO.layer_norm(h, gamma, beta, self._config.eps, r)
        elif r is not None:
            h = h + r
        if isinstance(h, tuple):
            h = h[0]
        if h.is_cuda and h.dtype == torch.bfloat16:
            logits = gemm(h, vocab_embedding)
        else:
            logits = torch.nn.functional.linear(h, vocab_embedding)
        if bias is not None:
            logits = logits + bias
        self._output = logits
        return logits
