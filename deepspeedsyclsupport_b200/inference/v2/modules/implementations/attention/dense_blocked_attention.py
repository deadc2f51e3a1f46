"""Attention against the blocked KV cache (parity: inference/v2/modules/implementations/attention/dense_blocked_attention.py:24-
169: linear_blocked_kv_rotary + blocked_flash).

Two launches per layer: `kv_rotary_append` rotates q and k in place and writes k / v rows to their cache blocks, then
`paged_attention` runs every token of the ragged batch (prefill chunks and decode steps alike) against its sequence's
blocks.  Both read the batch metadata the RaggedBatchWrapper already holds on the device, so `build_atoms` has nothing to add."""
import math
from typing import Any, Dict, Optional

import torch

from ......ops import fused_ops as FO
from .... import kernels as K
from ....inference_utils import DtypeEnum
from ...configs import DSSelfAttentionConfig, MaskingType, PositionalEmbeddingType
from ...interfaces import DSSelfAttentionBase, DSSelfAttentionRegistry


@DSSelfAttentionRegistry.register_module
class DSDenseBlockedAttention(DSSelfAttentionBase):

    @staticmethod
    def name() -> str:
        return "dense_blocked_attention"

    @staticmethod
    def supports_config(config: DSSelfAttentionConfig) -> bool:
        if config.input_dtype != config.output_dtype:
            return False
        if DtypeEnum(config.input_dtype) not in (DtypeEnum.fp16, DtypeEnum.bf16, DtypeEnum.fp32):
            return False
        if PositionalEmbeddingType(config.positional_embedding_type) not in (PositionalEmbeddingType.none, PositionalEmbeddingType.rotate_half):
            return False
        if MaskingType(config.masking_type) != MaskingType.causal:
            return False
        return config.n_heads_q % config.n_heads_kv == 0

    def __init__(self, config: DSSelfAttentionConfig, implementation_config: Dict[str, Any] = None) -> None:
        super().__init__(config, implementation_config)
        c = config
        self._rotary = PositionalEmbeddingType(c.positional_embedding_type) == PositionalEmbeddingType.rotate_half
        self._rot_dim, self._inv_freq, self._trained = 0, None, False
        if self._rotary:
            rc = c.positional_embedding_config
            self._rot_dim = c.head_size if rc is None or rc.rotate_dim is None else int(rc.rotate_dim)
            self._theta = 10000.0 if rc is None else rc.theta_base
            self._trained = bool(rc is not None and rc.use_trained_freqs)
        # scale_factor is the full softmax scale (the model passes 1/sqrt(d) or its own variant)
        self._scale = float(c.scale_factor) if c.scale_factor != 1.0 else 1.0 / math.sqrt(c.head_size)

    @property
    def kv_block_size(self) -> int:
        return 128

    @property
    def q_block_size(self) -> int:
        return 128

    def _freqs(self, device, inv_freqs):
        if self._trained:
            assert inv_freqs is not None, "use_trained_freqs needs inv_freqs"
            return inv_freqs
        if self._inv_freq is None or self._inv_freq.device != device:
            self._inv_freq = FO.rope_inv_freq(self._rot_dim, self._theta, device=device)
        return self._inv_freq

    def forward(self, q_k_v: torch.Tensor, kv_cache: torch.Tensor, batch, attention_mask: Optional[torch.Tensor] = None,
                attention_bias: Optional[torch.Tensor] = None, inv_freqs: Optional[torch.Tensor] = None) -> torch.Tensor:
        c = self._config
        tok_seq, tok_pos, bt = batch.tok_seq(), batch.tok_pos(), batch.block_table()
        freqs = self._freqs(q_k_v.device, inv_freqs) if self._rotary else None
        K.kv_rotary_append(q_k_v, kv_cache, tok_seq, tok_pos, bt, freqs, c.n_heads_q, c.n_heads_kv, c.head_size,
                           rot_dim=self._rot_dim if self._rotary else 0)
        return K.paged_attention(q_k_v, kv_cache, tok_seq, tok_pos, bt, c.n_heads_q, c.n_heads_kv, c.head_size, self._scale,
                                 decode=batch.current_tokens <= 2 * batch.current_sequences)
