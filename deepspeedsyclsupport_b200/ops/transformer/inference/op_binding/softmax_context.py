"""(parity: op_binding/softmax_context.py)  Attention core over a fused QKV tensor: split heads, (rotary by config), append to
the layer's KV cache, masked softmax(QK^T) V.  Returns (context [B, S, H*D], key, value).

Plain causal steps in bf16 (head size 64 / 128, no padding mask / ALiBi / local window) run on the KV arena + tcgen05 paged flash
kernel (kv_arena.py), as csrc/transformer/inference/csrc/pt_binding.cpp:434-558 runs on the context workspace; everything else
takes the general masked-softmax formulation below."""
import torch

from ..ds_attention import apply_rotary
from ..kv_arena import KVArena
from ..kv_arena import eligible as arena_eligible
from .base import BaseOp


class SoftmaxContextOp(BaseOp):

    def __init__(self, config):
        super().__init__(config)
        self._cache = {}
        self._arenas = {}

    def reset(self):
        self._cache = {}
        for a in self._arenas.values():
            a.reset()

    def forward(self, query_key_value: torch.Tensor, attn_mask: torch.Tensor, heads: int, num_kv: int, norm_factor: float, no_masking: bool,
                layer_id: int, num_layers: int, alibi: torch.Tensor):
        cfg = self.config
        B, S, _ = query_key_value.shape
        kv = num_kv if num_kv and num_kv > 0 else heads
        d = query_key_value.shape[-1] // (heads + 2 * kv)
        masked = cfg.triangular_masking and not no_masking
        plain = (attn_mask is None or attn_mask.numel() <= 1 or B == 1) and (alibi is None or alibi.numel() <= 1)
        if masked and plain and not cfg.local_attention and layer_id not in self._cache and arena_eligible(query_key_value, d, heads, kv, cfg, None, None, None):
            # tensor-core path: rotary + append into the layer's arena, paged flash attention (kv_arena.py)
            arena = self._arenas.get(layer_id)
            if arena is None:
                arena = self._arenas[layer_id] = KVArena(min_tokens=max(int(getattr(cfg, "max_out_tokens", 1024) or 1024), 128))
            if S > 1:
                arena.reset()
            qkv = query_key_value.contiguous()
            inv_freq, rot = None, 0
            if cfg.rotary_dim > 0 and cfg.rotate_half:
                rd = cfg.rotary_dim
                inv_freq = 1.0 / (cfg.rope_theta**(torch.arange(0, rd, 2, device=qkv.device, dtype=torch.float32) / rd))
                rot = rd
            elif cfg.rotary_dim > 0:
                qkv = qkv.clone()
                x = qkv.view(B, S, heads + 2 * kv, d)
                pos = torch.arange(arena.length, arena.length + S, device=qkv.device)
                x[:, :, :heads + kv] = apply_rotary(x[:, :, :heads + kv], pos, cfg.rotary_dim, cfg.rope_theta, False)
            elif qkv.data_ptr() == query_key_value.data_ptr():
                qkv = qkv.clone()  # the append kernel may rotate in place: never touch the caller's tensor
            ctx = arena.attend(qkv, heads, kv, d, norm_factor * norm_factor, inv_freq, rot)
            keys, values = arena.keys_values()
            return ctx, keys, values
        q, k, v = query_key_value.view(B, S, heads + 2 * kv, d).split([heads, kv, kv], 2)
        past_k, past_v = self._cache.get(layer_id, (None, None)) if S == 1 else (None, None)
        past = 0 if past_k is None else past_k.shape[2]
        if cfg.rotary_dim > 0:
            pos = torch.arange(past, past + S, device=q.device)
            q, k = apply_rotary(q, pos, cfg.rotary_dim, cfg.rope_theta, cfg.rotate_half), apply_rotary(k, pos, cfg.rotary_dim, cfg.rope_theta,
                                                                                                  cfg.rotate_half)
        k, v = k.transpose(1, 2), v.transpose(1, 2)
        if past_k is not None:
            k, v = torch.cat([past_k, k], 2), torch.cat([past_v, v], 2)
        self._cache[layer_id] = (k, v)
        T = k.shape[2]
        rep = heads // kv
        kk, vv = (k.repeat_interleave(rep, 1), v.repeat_interleave(rep, 1)) if rep > 1 else (k, v)
        bias = None
        if cfg.triangular_masking and not no_masking:
            i = torch.arange(past, past + S, device=q.device)[:, None]
            j = torch.arange(T, device=q.device)[None, :]
            allowed = j <= i
            if cfg.local_attention:
                allowed = allowed & (j > i - cfg.window_size)
            bias = torch.zeros(S, T, device=q.device).masked_fill(~allowed, float("-inf"))[None, None]
        if attn_mask is not None and attn_mask.numel() > 1:
            m = attn_mask.float()
            while m.dim() < 4:
                m = m[:, None]
            bias = m if bias is None else bias + m
        if alibi is not None and alibi.numel() > 1:
            a = alibi.reshape(B, heads, 1, -1)[..., :T].float()
            bias = a if bias is None else bias + a
        # norm_factor is applied to q and k each in the reference kernel: the score scale is its square
        ctx = torch.nn.functional.scaled_dot_product_attention(q.transpose(1, 2), kk, vv, attn_mask=None if bias is None else bias.to(q.dtype),
                                                               scale=norm_factor * norm_factor)
        return ctx.transpose(1, 2).reshape(B, S, heads * d), k, v
