"""Ragged-batch kernels as configured objects (parity: inference/v2/kernels/ragged_ops/{atom_builder, blocked_flash, embed,
linear_blocked_kv_rotary, logits_gather, moe_gather, moe_scatter, top_k_gating}).  Device code: csrc/ragged/ragged_ops.cu; the
functional forms live in kernels/__init__.py and have host restatements used by the CPU tests."""
from typing import Optional

import torch

from ... import kernels as K
from ...inference_utils import DtypeEnum
from ..ds_kernel import DSKernelBase

_FP = (DtypeEnum.fp16, DtypeEnum.bf16, DtypeEnum.fp32)


def _fp(dtype, who):
    if DtypeEnum(dtype) not in _FP:
        raise ValueError(f"Unsupported data type: {dtype} for {who}")
    return DtypeEnum(dtype).value


class AtomBuilder(DSKernelBase):
    """Work list of the blocked attention kernel: one atom per (sequence, q block): (q start, q length, KV blocks visible).
    Host-side (csrc/host/ragged_host.cpp through RaggedBatchWrapper.atoms); returns (atoms, number of atoms)."""

    def __init__(self) -> None:
        pass

    def __call__(self, atoms: Optional[torch.Tensor], ragged_batch, q_block_size: int, kv_block_size: int):
        built = ragged_batch.atoms(q_block_size, kv_block_size)
        if atoms is not None:
            n = min(atoms.shape[0], built.shape[0])
            atoms[:n].copy_(built[:n])
            return atoms, n
        return built, built.shape[0]


def _check_head_size(head_size, dtype):
    """The device kernels are instantiated for these head sizes; fp32 runs the host restatement, which takes any size."""
    if dtype != torch.float32 and head_size not in _KVKernel.supported_head_sizes:
        raise ValueError(f"Unsupported head size: {head_size}, supported_head_sizes are {_KVKernel.supported_head_sizes}")


class _KVKernel(DSKernelBase):
    supported_head_sizes = (64, 96, 128, 256)

    def __init__(self, head_size: int, n_q_heads: int, n_kv_heads: int, dtype) -> None:
        self.dtype = _fp(dtype, type(self).__name__)
        _check_head_size(head_size, self.dtype)
        if n_q_heads % n_kv_heads != 0:
            raise ValueError(f"n_q_heads ({n_q_heads}) must be a multiple of n_kv_heads ({n_kv_heads})")
        self.head_size, self.n_q_heads, self.n_kv_heads = head_size, n_q_heads, n_kv_heads

    def _run(self, kv_cache, qkv, ragged_batch, inv_freq, rot_dim):
        K.kv_rotary_append(qkv, kv_cache, ragged_batch.tok_seq(), ragged_batch.tok_pos(), ragged_batch.block_table(), inv_freq, self.n_q_heads,
                           self.n_kv_heads, self.head_size, rot_dim=rot_dim)


class LinearBlockedKVCopy(_KVKernel):
    """Append this batch's K / V rows to their cache blocks; no positional rotation."""

    def __call__(self, kv_cache: torch.Tensor, qkv: torch.Tensor, ragged_batch) -> None:
        self._run(kv_cache, qkv, ragged_batch, None, 0)


class BlockedRotaryEmbeddings(_KVKernel):
    """Rotate q and k in place with synthesised frequencies theta_base ** (-2i / rotary_dim), then append K / V to the cache."""

    def __init__(self, head_size: int, n_q_heads: int, n_kv_heads: int, dtype, rotary_dim: int, theta_base: float) -> None:
        super().__init__(head_size, n_q_heads, n_kv_heads, dtype)
        if rotary_dim > head_size or rotary_dim % 2:
            raise ValueError(f"rotary_dim must be even and <= head_size, got {rotary_dim}")
        self.rotary_dim, self.theta_base, self._freqs = rotary_dim, theta_base, None

    def __call__(self, kv_cache: torch.Tensor, qkv: torch.Tensor, ragged_batch) -> None:
        if self._freqs is None or self._freqs.device != qkv.device:
            from .....ops.fused_ops import rope_inv_freq
            self._freqs = rope_inv_freq(self.rotary_dim, self.theta_base, device=qkv.device)
        self._run(kv_cache, qkv, ragged_batch, self._freqs, self.rotary_dim)


class BlockedTrainedRotaryEmbeddings(_KVKernel):
    """Same with inverse frequencies taken from the checkpoint."""

    def __call__(self, kv_cache: torch.Tensor, qkv: torch.Tensor, ragged_batch, inverse_freqs: torch.Tensor) -> None:
        self._run(kv_cache, qkv, ragged_batch, inverse_freqs, 2 * inverse_freqs.numel())


class BlockedFlashAttn(DSKernelBase):
    """Causal attention of every token of the ragged batch against its sequence's cache blocks."""
    supported_head_sizes = _KVKernel.supported_head_sizes

    def __init__(self, head_size: int, dtype) -> None:
        self.dtype = _fp(dtype, "BlockedFlashAttn")
        _check_head_size(head_size, self.dtype)
        self.head_size = head_size

    def __call__(self, out: torch.Tensor, qkv: torch.Tensor, kv_cache: torch.Tensor, ragged_batch, n_q_heads: int, n_kv_heads: int,
                 softmax_scale: Optional[float] = None) -> torch.Tensor:
        out.copy_(K.paged_attention(qkv, kv_cache, ragged_batch.tok_seq(), ragged_batch.tok_pos(), ragged_batch.block_table(), n_q_heads,
                                    n_kv_heads, self.head_size, softmax_scale))
        return out


class RaggedEmbeddingKernel(DSKernelBase):

    def __init__(self, embed_dtype, token_dtype, embed_dim: int) -> None:
        self.dtype = _fp(embed_dtype, "RaggedEmbeddingKernel")
        if token_dtype not in (torch.int32, torch.int64):
            raise ValueError(f"Unsupported token dtype: {token_dtype}, supported are int32 and int64")
        if (embed_dim * torch.empty((), dtype=self.dtype).element_size()) % 16 != 0:
            raise ValueError(f"Embedding rows must be a multiple of 16 bytes, got dim {embed_dim} of {self.dtype}")
        self.embed_dim = embed_dim

    def __call__(self, embedded_tokens: torch.Tensor, ragged_wrapper, embedding_weight: torch.Tensor,
                 position_embed_weight: Optional[torch.Tensor] = None, position_embed_offset: int = 0) -> torch.Tensor:
        embedded_tokens.copy_(K.ragged_embed(ragged_wrapper.input_ids(), embedding_weight, position_embed_weight,
                                             ragged_wrapper.tok_pos() if position_embed_weight is not None else None, position_embed_offset))
        return embedded_tokens


class RaggedLogitsGather(DSKernelBase):

    def __init__(self, model_dim: int, fp_dtype):
        self.dtype = _fp(fp_dtype, "RaggedLogitsGather")
        if (model_dim * torch.empty((), dtype=self.dtype).element_size()) % 16 != 0:
            raise ValueError(f"Embedding dimension must be aligned to 16 bytes, got {model_dim}")

    def __call__(self, final_token_activations: torch.Tensor, all_activations: torch.Tensor, ragged_wrapper) -> torch.Tensor:
        final_token_activations.copy_(K.logits_gather(all_activations, ragged_wrapper.last_tokens()))
        return final_token_activations


class RaggedTopKGating(DSKernelBase):
    """softmax over experts in fp32, top-k, per-expert token counts and each token's offset inside its expert's range."""

    def __init__(self, logit_dtype) -> None:
        self.dtype = _fp(logit_dtype, "RaggedTopKGating")

    def __call__(self, expert_counts: torch.Tensor, scores: torch.Tensor, assignments: torch.Tensor, offsets: torch.Tensor, logits: torch.Tensor,
                 batch=None, normalize: bool = True):
        c, s, a, o = K.top_k_gating(logits.float(), scores.shape[1], normalize=normalize)
        expert_counts.copy_(c), scores.copy_(s), assignments.copy_(a), offsets.copy_(o)
        return expert_counts, scores, assignments, offsets


class MoEScatter(DSKernelBase):

    def __init__(self, dtype, channels: int) -> None:
        self.dtype = _fp(dtype, "MoEScatter")
        if (channels * torch.empty((), dtype=self.dtype).element_size()) % 16 != 0:
            raise ValueError(f"channels must be aligned to 16 bytes, got {channels}")

    def __call__(self, moe_input: torch.Tensor, expert_cumsum: torch.Tensor, mapped_slots: torch.Tensor, activations: torch.Tensor,
                 expert_counts: torch.Tensor, assignments: torch.Tensor, offsets: torch.Tensor):
        m, c, s = K.moe_scatter(activations, expert_counts, assignments, offsets)
        moe_input.copy_(m), expert_cumsum.copy_(c), mapped_slots.copy_(s)
        return moe_input, expert_cumsum, mapped_slots


class MoEGather(DSKernelBase):

    def __init__(self, dtype, channels: int, normalize_scores: bool = False) -> None:
        self.dtype = _fp(dtype, "MoEGather")
        if (channels * torch.empty((), dtype=self.dtype).element_size()) % 16 != 0:
            raise ValueError(f"channels must be aligned to 16 bytes, got {channels}")
        self.normalize_scores = normalize_scores

    def __call__(self, layer_output: torch.Tensor, moe_output: torch.Tensor, scores: torch.Tensor, mapped_slots: torch.Tensor,
                 expert_counts: Optional[torch.Tensor] = None) -> torch.Tensor:
        if self.normalize_scores:
            scores = scores / scores.sum(-1, keepdim=True)
        layer_output.copy_(K.moe_gather(moe_output, scores, mapped_slots))
        return layer_output
