"""GPT-2 (small/medium/large/xl shapes) on the framework's kernels: LayerNorm, fused QKV GEMM, flash attention,
bias + GELU.  BASELINE.json config #1 (GPT-2 small, ZeRO-1, gloo world 2) runs this model on the host path."""
from dataclasses import dataclass

import torch
import torch.nn as nn

from ..ops import fused_ops as FO
from ..ops.attention import attention
from ..ops.gemm import linear


@dataclass
class GPT2Config:
    vocab_size: int = 50257
    n_positions: int = 1024
    n_embd: int = 768
    n_layer: int = 12
    n_head: int = 12
    layer_norm_epsilon: float = 1e-5
    initializer_range: float = 0.02
    activation_checkpointing: bool = False


GPT2_SHAPES = {"small": dict(n_embd=768, n_layer=12, n_head=12), "medium": dict(n_embd=1024, n_layer=24, n_head=16),
               "large": dict(n_embd=1280, n_layer=36, n_head=20), "xl": dict(n_embd=1600, n_layer=48, n_head=25),
               "tiny": dict(n_embd=64, n_layer=2, n_head=2, vocab_size=256, n_positions=64)}


def gpt2_config(name="small", **over):
    kw = dict(GPT2_SHAPES[name])
    kw.update(over)
    return GPT2Config(**kw)


class LayerNorm(nn.Module):

    def __init__(self, d, eps):
        super().__init__()
        self.weight = nn.Parameter(torch.ones(d))
        self.bias = nn.Parameter(torch.zeros(d))
        self.eps = eps

    def forward(self, x):
        return FO.layer_norm(x, self.weight, self.bias, self.eps)


class GPT2Block(nn.Module):

    def __init__(self, cfg):
        super().__init__()
        d = cfg.n_embd
        self.cfg = cfg
        self.ln_1 = LayerNorm(d, cfg.layer_norm_epsilon)
        self.c_attn_w = nn.Parameter(torch.empty(3 * d, d))
        self.c_attn_b = nn.Parameter(torch.zeros(3 * d))
        self.c_proj_w = nn.Parameter(torch.empty(d, d))
        self.c_proj_b = nn.Parameter(torch.zeros(d))
        self.ln_2 = LayerNorm(d, cfg.layer_norm_epsilon)
        self.fc_w = nn.Parameter(torch.empty(4 * d, d))
        self.fc_b = nn.Parameter(torch.zeros(4 * d))
        self.proj_w = nn.Parameter(torch.empty(d, 4 * d))
        self.proj_b = nn.Parameter(torch.zeros(d))
        for w in (self.c_attn_w, self.c_proj_w, self.fc_w, self.proj_w):
            nn.init.normal_(w, std=cfg.initializer_range)

    def forward(self, x):
        B, S, d = x.shape
        h = self.ln_1(x)
        qkv = linear(h, self.c_attn_w) + self.c_attn_b
        a = attention(qkv.view(B * S, -1).contiguous(), B, S, self.cfg.n_head, self.cfg.n_head, d // self.cfg.n_head, causal=True)
        x = x + linear(a.view(B, S, d), self.c_proj_w) + self.c_proj_b
        h = self.ln_2(x)
        h = torch.nn.functional.gelu(linear(h, self.fc_w) + self.fc_b, approximate="tanh")
        return x + linear(h, self.proj_w) + self.proj_b


class GPT2LMHeadModel(nn.Module):

    def __init__(self, cfg: GPT2Config):
        super().__init__()
        self.cfg = cfg
        self.wte = nn.Embedding(cfg.vocab_size, cfg.n_embd)
        self.wpe = nn.Embedding(cfg.n_positions, cfg.n_embd)
        nn.init.normal_(self.wte.weight, std=cfg.initializer_range)
        nn.init.normal_(self.wpe.weight, std=cfg.initializer_range)
        self.h = nn.ModuleList([GPT2Block(cfg) for _ in range(cfg.n_layer)])
        self.ln_f = LayerNorm(cfg.n_embd, cfg.layer_norm_epsilon)

    def forward(self, input_ids, labels=None):
        B, S = input_ids.shape
        pos = torch.arange(S, device=input_ids.device)
        x = self.wte(input_ids) + self.wpe(pos)[None]
        for blk in self.h:
            if self.cfg.activation_checkpointing and self.training:
                x = torch.utils.checkpoint.checkpoint(blk, x, use_reentrant=False)
            else:
                x = blk(x)
        x = self.ln_f(x)
        logits = torch.nn.functional.linear(x, self.wte.weight)  # tied embedding
        if labels is None:
            return logits
        return FO.cross_entropy(logits.view(-1, logits.shape[-1]), labels.view(-1))
