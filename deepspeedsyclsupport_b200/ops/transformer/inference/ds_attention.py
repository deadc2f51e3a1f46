"""Self-attention half of the inference block with an in-module KV cache (reference: ops/transformer/inference/ds_attention.py
DeepSpeedSelfAttention over csrc/transformer/inference `qkv_gemm` / `softmax_context` / `vector_matmul`).

Parameters keep the reference names (attn_qkvw / attn_qkvb / attn_ow / attn_ob) but are stored [out, in] -- the native layout of
the tcgen05 GEMM (`transposed_mode` of the reference).  QKV is one fused projection; GQA through `config.num_kv`; rotary
(`rotary_dim`, rotate_half or rotate_every_two) is applied to the new tokens before they are appended to the cache.

Plain causal steps (bf16, head size 64 / 128, no padding mask, no ALiBi) run on the layer's KV arena (kv_arena.py): rotary+append
kernel, then the tcgen05 paged flash kernel for prompt and decode alike.  Masked / biased steps keep a dense cache and the
general masked-softmax formulation."""
import math

import torch
from torch import nn

from ....accelerator import get_accelerator
from ... import gemm as G
from .kv_arena import KVArena
from .kv_arena import alibi_slopes_of
from .kv_arena import eligible as arena_eligible


def _linear(x, w, b):
    if x.is_cuda and x.dtype == torch.bfloat16 and w.dtype == torch.bfloat16:
        y = G.gemm(x.reshape(-1, x.shape[-1]), w).view(*x.shape[:-1], w.shape[0])
        return y if b is None else y + b
    return torch.nn.functional.linear(x, w, b)


def apply_rotary(x, positions, rotary_dim, theta, rotate_half):
    """x: [B, S, H, D]; rotates the first rotary_dim features of every head at the given absolute positions [S]."""
    d = rotary_dim
    inv = 1.0 / (theta**(torch.arange(0, d, 2, device=x.device, dtype=torch.float32) / d))
    ang = positions.to(torch.float32)[:, None] * inv[None, :]  # [S, d/2]
    cos, sin = ang.cos()[None, :, None, :], ang.sin()[None, :, None, :]
    xr, rest = x[..., :d].float(), x[..., d:]
    if rotate_half:
        a, b = xr[..., :d // 2], xr[..., d // 2:]
        out = torch.cat([a * cos - b * sin, b * cos + a * sin], -1)
    else:  # rotate_every_two (GPT-J): pairs (2i, 2i+1)
        a, b = xr[..., 0::2], xr[..., 1::2]
        out = torch.stack([a * cos - b * sin, b * cos + a * sin], -1).flatten(-2)
    return torch.cat([out.to(x.dtype), rest], -1)


class DeepSpeedSelfAttention(nn.Module):
    num_layers = 0

    def __init__(self, config, mp_group=None, q_scales=None, q_groups=1, merge_count=1):
        super().__init__()
        self.config = config
        self.mp_group = mp_group
        dt = torch.half if config.dtype == torch.int8 else config.dtype
        dev = get_accelerator().current_device_name()
        h, mp = config.hidden_size, config.mp_size
        self.heads = config.heads // mp
        self.kv_heads = (config.num_kv if config.num_kv > 0 else config.heads) // mp
        self.head_dim = h // config.heads
        qkv_out = (self.heads + 2 * self.kv_heads) * self.head_dim
        mk = lambda *s: nn.Parameter(torch.empty(*s, dtype=dt, device=dev), requires_grad=False)
        if config.set_empty_params:
            self.attn_qkvw = self.attn_qkvb = self.attn_ow = self.attn_ob = None
        else:
            self.attn_qkvw, self.attn_qkvb = mk(qkv_out, h), mk(qkv_out)
            self.attn_ow, self.attn_ob = mk(h, self.heads * self.head_dim), mk(h)
        self.layer_id = DeepSpeedSelfAttention.num_layers
        DeepSpeedSelfAttention.num_layers += 1
        self.scale = 1.0 / math.sqrt(self.head_dim) if config.scale_attention else 1.0
        if config.scale_attn_by_inverse_layer_idx:
            self.scale /= float(self.layer_id + 1)
        self.k_cache = self.v_cache = None  # [B, heads_kv, T, D] -- masked / biased path only
        self.arena = KVArena(min_tokens=max(int(getattr(config, "max_out_tokens", 1024) or 1024), 128))
        self._inv_freq = None
        self._alibi_slopes = None  # per-head ALiBi slopes read off the first alibi tensor seen (BLOOM)

    def reset_cache(self):
        self.k_cache = self.v_cache = None
        self.arena.reset()

    def _forward_arena(self, qkv, B, S, get_present, alibi=None):
        """Plain causal attention on the tensor-core kernels: rotary + append into the layer's KV arena, paged flash attention."""
        cfg = self.config
        inv_freq, rot = None, 0
        if cfg.rotary_dim > 0:
            if cfg.rotate_half:
                if self._inv_freq is None or self._inv_freq.device != qkv.device:
                    d = cfg.rotary_dim
                    self._inv_freq = 1.0 / (cfg.rope_theta**(torch.arange(0, d, 2, device=qkv.device, dtype=torch.float32) / d))
                inv_freq, rot = self._inv_freq, cfg.rotary_dim
            else:  # rotate_every_two pairing (GPT-J): rotated here, the kernel only appends
                pos = torch.arange(self.arena.length, self.arena.length + S, device=qkv.device)
                x = qkv.view(B, S, self.heads + 2 * self.kv_heads, self.head_dim)
                n = self.heads + self.kv_heads
                x[:, :, :n] = apply_rotary(x[:, :, :n], pos, cfg.rotary_dim, cfg.rope_theta, False)
        slopes = None
        if alibi is not None:
            if self._alibi_slopes is None:
                self._alibi_slopes = alibi_slopes_of(alibi, B, self.heads)
            slopes = self._alibi_slopes
        window = int(cfg.window_size) if getattr(cfg, "local_attention", False) else 0
        ctx = self.arena.attend(qkv, self.heads, self.kv_heads, self.head_dim, self.scale, inv_freq, rot, alibi_slopes=slopes, window=window)
        out = _linear(ctx, self.attn_ow, None)
        if self.mp_group is not None and cfg.mp_size > 1:
            from .... import comm as dist
            dist.all_reduce(out, group=self.mp_group)
        if self.attn_ob is not None:
            out = out + self.attn_ob
        if get_present:
            k, v = self.arena.keys_values()
        else:
            x = qkv.view(B, S, self.heads + 2 * self.kv_heads, self.head_dim)
            k, v = x[:, :, self.heads:self.heads + self.kv_heads].transpose(1, 2), x[:, :, self.heads + self.kv_heads:].transpose(1, 2)
        return out, k, v

    def forward(self, input, input_mask=None, head_mask=None, layer_past=None, get_present=False, encoder_hidden_states=None,
                encoder_attention_mask=None, output_attentions=False, norm_w=None, norm_b=None, alibi=None):
        """input: [B, S, hidden] already normalised by the caller (pre-LN) -- returns (attn_out [B,S,hidden], key, value)."""
        B, S, _ = input.shape
        qkv = _linear(input, self.attn_qkvw, self.attn_qkvb)
        if layer_past is None and S > 1:
            self.reset_cache()  # a new prompt
        if self.k_cache is None and arena_eligible(qkv, self.head_dim, self.heads, self.kv_heads, self.config, input_mask, alibi, layer_past):
            return self._forward_arena(qkv, B, S, get_present, alibi)
        if self.k_cache is None and self.arena.length > 0:  # a decode step the arena kernels cannot take: continue densely
            self.k_cache, self.v_cache = (t.contiguous() for t in self.arena.keys_values())
            self.arena.reset()
        q, k, v = qkv.view(B, S, self.heads + 2 * self.kv_heads, self.head_dim).split([self.heads, self.kv_heads, self.kv_heads], 2)
        if layer_past is not None:
            self.k_cache, self.v_cache = layer_past
        past = 0 if self.k_cache is None else self.k_cache.shape[2]
        cfg = self.config
        if cfg.rotary_dim > 0:
            pos = torch.arange(past, past + S, device=input.device)
            q = apply_rotary(q, pos, cfg.rotary_dim, cfg.rope_theta, cfg.rotate_half)
            k = apply_rotary(k, pos, cfg.rotary_dim, cfg.rope_theta, cfg.rotate_half)
        k, v = k.transpose(1, 2), v.transpose(1, 2)
        if self.k_cache is not None:
            k, v = torch.cat([self.k_cache, k], 2), torch.cat([self.v_cache, v], 2)
        self.k_cache, self.v_cache = k, v
        T = k.shape[2]
        rep = self.heads // self.kv_heads
        kk, vv = (k.repeat_interleave(rep, 1), v.repeat_interleave(rep, 1)) if rep > 1 else (k, v)
        bias = None
        if cfg.triangular_masking:
            i = torch.arange(past, past + S, device=input.device)[:, None]
            j = torch.arange(T, device=input.device)[None, :]
            allowed = j <= i
            if cfg.local_attention:
                allowed = allowed & (j > i - cfg.window_size)
            bias = torch.zeros(S, T, device=input.device, dtype=torch.float32).masked_fill(~allowed, float("-inf"))[None, None]
        if input_mask is not None:
            m = input_mask.to(torch.float32)
            while m.dim() < 4:
                m = m[:, None]
            bias = m if bias is None else bias + m
        if alibi is not None:
            a = alibi.view(B, self.heads, 1, -1)[..., :T].to(torch.float32)
            bias = a if bias is None else bias + a
        ctx = torch.nn.functional.scaled_dot_product_attention(q.transpose(1, 2), kk, vv, attn_mask=None if bias is None else bias.to(q.dtype),
                                                               scale=self.scale)
        out = _linear(ctx.transpose(1, 2).reshape(B, S, -1), self.attn_ow, None)
        if self.mp_group is not None and cfg.mp_size > 1:
            from .... import comm as dist
            dist.all_reduce(out, group=self.mp_group)
        if self.attn_ob is not None:
            out = out + self.attn_ob
        return out, k, v
