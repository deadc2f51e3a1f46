"""Llama-2 family (7B / 13B / 70B shapes) built on the framework's sm_100a kernels.

Per decoder layer: fused pre-RMSNorm(+residual add) -> one QKV GEMM (tcgen05) -> in-place RoPE on the packed q,k heads
-> flash attention on the packed projection -> output GEMM -> fused RMSNorm(+residual) -> one gate|up GEMM -> SiGLU
-> down GEMM.  The residual stream is carried as (x, residual) so every residual add is folded into the next norm
kernel.  Weight gradients are accumulated by the wgrad GEMM epilogue straight into the ZeRO gradient buffers.
The reference ships no model definitions (users bring HF modules); this module is the flagship workload of
BASELINE.json (Llama-2-7B ZeRO-3) and the source of the AutoTP / FastGen Llama containers.
"""
import math
from dataclasses import dataclass

import torch
import torch.nn as nn

from ..ops import fused_ops as FO
from ..ops.attention import attention
from ..ops import recompute as RC
from ..ops.gemm import linear, linear_rope


@dataclass
class LlamaConfig:
    vocab_size: int = 32000
    hidden_size: int = 4096
    intermediate_size: int = 11008
    num_hidden_layers: int = 32
    num_attention_heads: int = 32
    num_key_value_heads: int = 32
    max_position_embeddings: int = 4096
    rms_norm_eps: float = 1e-5
    rope_theta: float = 10000.0
    tie_word_embeddings: bool = False
    activation_checkpointing: bool = False
    # what a checkpointed layer keeps instead of recomputing (ops/recompute.py): full | attention | gemm | auto
    checkpoint_policy: str = "auto"
    save_attention_in_checkpoint: bool = True  # legacy switch: False forces checkpoint_policy = "full"
    initializer_range: float = 0.02
    # DeepSpeed-Ulysses: every rank of the group sees S / P consecutive tokens of each sequence (LlamaForCausalLM.set_sequence_parallel)
    sequence_parallel_group: object = None
    sequence_parallel_size: int = 1
    sequence_parallel_rank: int = 0

    @property
    def head_dim(self):
        return self.hidden_size // self.num_attention_heads


LLAMA2_7B = dict(hidden_size=4096, intermediate_size=11008, num_hidden_layers=32, num_attention_heads=32, num_key_value_heads=32)
LLAMA2_13B = dict(hidden_size=5120, intermediate_size=13824, num_hidden_layers=40, num_attention_heads=40, num_key_value_heads=40)
LLAMA2_70B = dict(hidden_size=8192, intermediate_size=28672, num_hidden_layers=80, num_attention_heads=64, num_key_value_heads=8)
TINY = dict(vocab_size=512, hidden_size=128, intermediate_size=384, num_hidden_layers=2, num_attention_heads=2, num_key_value_heads=2,
            max_position_embeddings=256)


def llama_config(name="7b", **overrides):
    base = {"7b": LLAMA2_7B, "13b": LLAMA2_13B, "70b": LLAMA2_70B, "tiny": TINY}[name.lower()]
    kw = dict(base)
    kw.update(overrides)
    return LlamaConfig(**kw)


class RMSNorm(nn.Module):

    def __init__(self, hidden, eps):
        super().__init__()
        self.weight = nn.Parameter(torch.ones(hidden))
        self.eps = eps

    def forward(self, x, residual=None):
        return FO.rms_norm(x, self.weight, self.eps, residual)


class LlamaDecoderLayer(nn.Module):

    def __init__(self, cfg: LlamaConfig, layer_idx=0):
        super().__init__()
        self.cfg = cfg
        h, d = cfg.hidden_size, cfg.head_dim
        self.hq, self.hkv, self.d = cfg.num_attention_heads, cfg.num_key_value_heads, d
        self.input_layernorm = RMSNorm(h, cfg.rms_norm_eps)
        self.qkv_proj = nn.Parameter(torch.empty((self.hq + 2 * self.hkv) * d, h))
        self.o_proj = nn.Parameter(torch.empty(h, self.hq * d))
        self.post_attention_layernorm = RMSNorm(h, cfg.rms_norm_eps)
        self.gate_up_proj = nn.Parameter(torch.empty(2 * cfg.intermediate_size, h))
        self.down_proj = nn.Parameter(torch.empty(h, cfg.intermediate_size))
        for w in (self.qkv_proj, self.o_proj, self.gate_up_proj, self.down_proj):
            nn.init.normal_(w, mean=0.0, std=cfg.initializer_range)
        self.register_buffer("inv_freq", FO.rope_inv_freq(d, cfg.rope_theta), persistent=False)

    def forward(self, x, residual, positions=None):
        """(x, residual) -> (mlp_out, new_residual); the true hidden state is x + residual."""
        B, S, _ = x.shape
        if residual is None:
            residual = x
            hn = self.input_layernorm(x)
        else:
            hn, residual = self.input_layernorm(x, residual)
        qkv = linear_rope(hn, self.qkv_proj, self.inv_freq, S, self.hq + self.hkv, self.d, positions, tag=RC.QKV)
        sp = self.cfg.sequence_parallel_group
        if sp is not None and self.cfg.sequence_parallel_size > 1:
            # DeepSpeed-Ulysses: S is this rank's slice of the sequence; re-shard to (whole sequence, this rank's heads) around
            # the local attention (sequence/fused.py: one-sided peer kernels that also do the layout change)
            from ..sequence.fused import attn_to_sequence, qkv_to_heads
            P = self.cfg.sequence_parallel_size
            qkv_h = qkv_to_heads(qkv.view(B * S, -1), B, S, self.hq, self.hkv, self.d, sp)
            attn_h = attention(qkv_h, B, S * P, self.hq // P, self.hkv // P, self.d, causal=True)
            attn = attn_to_sequence(attn_h.view(B * S * P, -1), B, S, self.hq, self.d, sp)
        else:
            attn = attention(qkv, B, S, self.hq, self.hkv, self.d, causal=True)
        attn_out = linear(attn.view(B, S, -1), self.o_proj, tag=RC.O_PROJ)
        hn, residual = self.post_attention_layernorm(attn_out, residual)
        act = FO.gated_act(linear(hn, self.gate_up_proj, tag=RC.GATE_UP), "silu")
        # the down projection's output is the layer's result: no backward formula reads it, so recomputation skips it
        return linear(act, self.down_proj, unused_in_recompute=True), residual


class LlamaModel(nn.Module):

    def __init__(self, cfg: LlamaConfig):
        super().__init__()
        self.cfg = cfg
        self.embed_tokens = nn.Embedding(cfg.vocab_size, cfg.hidden_size)
        nn.init.normal_(self.embed_tokens.weight, mean=0.0, std=cfg.initializer_range)
        self.layers = nn.ModuleList([LlamaDecoderLayer(cfg, i) for i in range(cfg.num_hidden_layers)])
        self.norm = RMSNorm(cfg.hidden_size, cfg.rms_norm_eps)

    def _checkpoint_policy(self, x):
        """Resolved once per (tokens) shape.  'auto': no checkpointing at all when every saved activation fits in 60 % of
        the free memory, else keep the GEMM outputs when they fit in half of it, else keep the attention core only."""
        cfg = self.cfg
        if not cfg.save_attention_in_checkpoint:
            return "full"
        tokens = x.shape[0] * x.shape[1]
        cached = getattr(self, "_policy_cache", None)
        if cached is None or cached[0] != tokens:
            d, h, inter = cfg.head_dim, cfg.hidden_size, cfg.intermediate_size
            qkv = (cfg.num_attention_heads + 2 * cfg.num_key_value_heads) * d
            per_token = 2 * (cfg.num_attention_heads * d + h + qkv + 2 * inter)  # bf16: attn out + o_proj out + qkv + gate_up
            kept = per_token * tokens * cfg.num_hidden_layers
            # everything autograd saves without checkpointing: 2 residuals, 2 norm outputs, qkv, attn out, gate_up, swiglu out
            everything = 2 * (4 * h + qkv + cfg.num_attention_heads * d + 3 * inter) * tokens * cfg.num_hidden_layers
            self._policy_cache = (tokens, RC.resolve_policy(cfg.checkpoint_policy, kept, everything))
        return self._policy_cache[1]

    def forward(self, input_ids, positions=None):
        x = self.embed_tokens(input_ids)
        if positions is None and self.cfg.sequence_parallel_size > 1:
            # this rank holds tokens [rank * S_local, (rank + 1) * S_local) of every sequence: RoPE needs their true positions
            B, S_local = input_ids.shape
            positions = (self.cfg.sequence_parallel_rank * S_local + torch.arange(S_local, device=input_ids.device)).repeat(B)
        residual = None
        for layer in self.layers:
            policy = self._checkpoint_policy(x) if (self.cfg.activation_checkpointing and self.training) else "none"
            if policy != "none":
                # per-layer checkpoint; the attention core's (out, lse) are kept and replayed during recomputation
                kw = dict(use_reentrant=False)
                if policy != "full":
                    kw["context_fn"] = lambda policy=policy: RC.contexts(policy)
                if residual is None:
                    x, residual = torch.utils.checkpoint.checkpoint(lambda a, l=layer: l(a, None, positions), x, **kw)
                else:
                    x, residual = torch.utils.checkpoint.checkpoint(lambda a, r, l=layer: l(a, r, positions), x, residual, **kw)
            else:
                x, residual = layer(x, residual, positions)
        if residual is None:
            return self.norm(x)
        out, _ = self.norm(x, residual)
        return out


class LlamaForCausalLM(nn.Module):
    """forward(input_ids, labels=None) -> loss (if labels) or logits.  Labels are next-token targets aligned with
    input positions (already shifted by the data pipeline); ignore_index = -100."""

    def __init__(self, cfg: LlamaConfig):
        super().__init__()
        self.cfg = cfg
        self.model = LlamaModel(cfg)
        if cfg.tie_word_embeddings:
            self.lm_head = None
        else:
            self.lm_head = nn.Parameter(torch.empty(cfg.vocab_size, cfg.hidden_size))
            nn.init.normal_(self.lm_head, mean=0.0, std=cfg.initializer_range)

    def set_sequence_parallel(self, group):
        """Turn on DeepSpeed-Ulysses over `group` (None = off): inputs are then the rank's S / P slice of every sequence, heads
        must divide by the group size, and the loss is the mean over the local tokens (average it over the group)."""
        from .. import comm as dist
        cfg = self.cfg
        cfg.sequence_parallel_group = group
        cfg.sequence_parallel_size = dist.get_world_size(group) if group is not None else 1
        cfg.sequence_parallel_rank = dist.get_rank(group) if group is not None else 0
        P = cfg.sequence_parallel_size
        assert cfg.num_attention_heads % P == 0 and cfg.num_key_value_heads % P == 0, "heads must split over the sequence-parallel group"
        return self

    def head_weight(self):
        return self.model.embed_tokens.weight if self.lm_head is None else self.lm_head

    def forward(self, input_ids, labels=None, positions=None):
        h = self.model(input_ids, positions)
        logits = linear(h, self.head_weight())
        if labels is None:
            return logits
        return FO.cross_entropy(logits.view(-1, logits.shape[-1]), labels.view(-1))

    def num_parameters(self):
        return sum(p.numel() for p in self.parameters())

    def flops_per_token(self, seq_len, with_recompute=False):
        """Training FLOPs per token (fwd+bwd = 3x fwd, +1 fwd with full recompute), attention included (causal)."""
        c = self.cfg
        d = c.head_dim
        per_layer = 2 * c.hidden_size * ((c.num_attention_heads + 2 * c.num_key_value_heads) * d + c.num_attention_heads * d
                                         + 3 * c.intermediate_size)
        attn = 2 * 2 * c.num_attention_heads * d * seq_len / 2  # QK^T and PV, causal half
        fwd = c.num_hidden_layers * (per_layer + attn) + 2 * c.hidden_size * c.vocab_size
        return fwd * (4 if with_recompute else 3)
