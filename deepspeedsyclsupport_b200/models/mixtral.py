"""Mixtral-style sparse MoE decoder (8x7B shapes) on the framework's MoE layer with expert parallelism.
Attention block is the Llama block; the MLP is `deepspeed.moe.layer.MoE` with SiGLU experts (top-2 gating)."""
from dataclasses import dataclass

import torch
import torch.nn as nn
import torch.utils.checkpoint

from ..moe.layer import MoE
from ..ops import fused_ops as FO
from ..ops.attention import attention
from ..ops.gemm import linear, linear_rope
from .llama import RMSNorm


@dataclass
class MixtralConfig:
    vocab_size: int = 32000
    hidden_size: int = 4096
    intermediate_size: int = 14336
    num_hidden_layers: int = 32
    num_attention_heads: int = 32
    num_key_value_heads: int = 8
    num_local_experts: int = 8
    num_experts_per_tok: int = 2
    rms_norm_eps: float = 1e-5
    rope_theta: float = 1e6
    ep_size: int = 1
    capacity_factor: float = 1.25
    router_aux_loss_coef: float = 0.02
    initializer_range: float = 0.02
    activation_checkpointing: bool = False

    @property
    def head_dim(self):
        return self.hidden_size // self.num_attention_heads


MIXTRAL_8X7B = dict()
MIXTRAL_TINY = dict(vocab_size=512, hidden_size=128, intermediate_size=256, num_hidden_layers=2, num_attention_heads=2,
                    num_key_value_heads=2, num_local_experts=4)


def mixtral_config(name="8x7b", **over):
    kw = dict(MIXTRAL_TINY if name == "tiny" else MIXTRAL_8X7B)
    kw.update(over)
    return MixtralConfig(**kw)


class SiGLUExpert(nn.Module):

    def __init__(self, hidden, inter, std=0.02):
        super().__init__()
        self.gate_up_proj = nn.Parameter(torch.empty(2 * inter, hidden))
        self.down_proj = nn.Parameter(torch.empty(hidden, inter))
        nn.init.normal_(self.gate_up_proj, std=std)
        nn.init.normal_(self.down_proj, std=std)

    def forward(self, x):
        shp = x.shape
        h = FO.gated_act(linear(x.reshape(-1, shp[-1]), self.gate_up_proj), "silu")
        return linear(h, self.down_proj).view(shp)


class MixtralDecoderLayer(nn.Module):

    def __init__(self, cfg: MixtralConfig):
        super().__init__()
        self.cfg = cfg
        h, d = cfg.hidden_size, cfg.head_dim
        self.hq, self.hkv, self.d = cfg.num_attention_heads, cfg.num_key_value_heads, d
        self.input_layernorm = RMSNorm(h, cfg.rms_norm_eps)
        self.qkv_proj = nn.Parameter(torch.empty((self.hq + 2 * self.hkv) * d, h))
        self.o_proj = nn.Parameter(torch.empty(h, self.hq * d))
        nn.init.normal_(self.qkv_proj, std=cfg.initializer_range)
        nn.init.normal_(self.o_proj, std=cfg.initializer_range)
        self.post_attention_layernorm = RMSNorm(h, cfg.rms_norm_eps)
        self.block_sparse_moe = MoE(hidden_size=h, expert=SiGLUExpert(h, cfg.intermediate_size, cfg.initializer_range),
                                    num_experts=cfg.num_local_experts, ep_size=cfg.ep_size, k=cfg.num_experts_per_tok,
                                    capacity_factor=cfg.capacity_factor, eval_capacity_factor=cfg.capacity_factor, min_capacity=4,
                                    use_rts=False)
        self.register_buffer("inv_freq", FO.rope_inv_freq(d, cfg.rope_theta), persistent=False)

    def forward(self, x):
        B, S, _ = x.shape
        hn = self.input_layernorm(x)
        qkv = linear_rope(hn, self.qkv_proj, self.inv_freq, S, self.hq + self.hkv, self.d)
        a = attention(qkv, B, S, self.hq, self.hkv, self.d, causal=True)
        x = x + linear(a.view(B, S, -1), self.o_proj)
        hn = self.post_attention_layernorm(x)
        moe_out, l_aux, _ = self.block_sparse_moe(hn)
        return x + moe_out, l_aux


class MixtralForCausalLM(nn.Module):

    def __init__(self, cfg: MixtralConfig):
        super().__init__()
        self.cfg = cfg
        self.embed_tokens = nn.Embedding(cfg.vocab_size, cfg.hidden_size)
        nn.init.normal_(self.embed_tokens.weight, std=cfg.initializer_range)
        self.layers = nn.ModuleList([MixtralDecoderLayer(cfg) for _ in range(cfg.num_hidden_layers)])
        self.norm = RMSNorm(cfg.hidden_size, cfg.rms_norm_eps)
        self.lm_head = nn.Parameter(torch.empty(cfg.vocab_size, cfg.hidden_size))
        nn.init.normal_(self.lm_head, std=cfg.initializer_range)

    def forward(self, input_ids, labels=None):
        x = self.embed_tokens(input_ids)
        aux = 0.0
        ckpt = self.cfg.activation_checkpointing and self.training and torch.is_grad_enabled()
        for layer in self.layers:
            if ckpt:  # per-layer recompute: the dispatched expert activations (capacity x 2 * intermediate per expert) are the bulk
                x, l_aux = torch.utils.checkpoint.checkpoint(layer, x, use_reentrant=False)
            else:
                x, l_aux = layer(x)
            aux = aux + l_aux
        logits = linear(self.norm(x), self.lm_head)
        if labels is None:
            return logits
        return FO.cross_entropy(logits.view(-1, logits.shape[-1]), labels.view(-1)) + self.cfg.router_aux_loss_coef * aux
