"""KV arena of the kernel-injection (v1) attention: the role of the reference's inference-context workspace
(csrc/transformer/inference/includes/inference_context.h:95-170 sizes one buffer for all layers' keys and values;
csrc/transformer/inference/csrc/pt_binding.cpp:434-558 `softmax_context` appends the new K/V rows to it and attends over it).

Here each attention layer owns one arena in the FastGen cache layout [blocks, 2, 128, kv_heads, head_dim]; sequence `b` of the
batch owns blocks [b * per_seq, (b + 1) * per_seq).  One call

    rotary (rotate-half form) + append of the new rows        csrc/ragged/ragged_ops.cu  kv_rotary_append_kernel
    causal attention of the new rows over the whole history   csrc/attn/paged_flash.cu   tcgen05 paged flash kernel

so prompt processing and single-token decode run on the same tensor-core kernel as FastGen, nothing is concatenated, and GQA
heads are packed into the MMA tile instead of being repeated.  The arena grows geometrically when a generation outruns it."""
import os

import torch

BLOCK = 128

_index_cache = {}


def _token_index(B, S, past, device):
    """(tok_seq, tok_pos) of a dense [B, S] step starting at position `past`; shared by all layers of a forward."""
    key = (B, S, past, device)
    hit = _index_cache.get(key)
    if hit is None:
        if len(_index_cache) > 8:
            _index_cache.clear()
        seq = torch.arange(B, device=device, dtype=torch.int32).repeat_interleave(S)
        pos = (torch.arange(S, device=device, dtype=torch.int32) + past).repeat(B)
        hit = _index_cache[key] = (seq, pos)
    return hit


def alibi_slopes_of(alibi, batch, heads):
    """Per-head slopes of a BLOOM-style ALiBi tensor ([batch * heads, 1, T] or [batch, heads, 1, T] = slope * key position for
    unpadded sequences), or None when they cannot be read off it (fewer than two key positions)."""
    al = alibi.reshape(batch, heads, -1)
    if al.shape[-1] < 2:
        return None
    return (al[0, :, 1] - al[0, :, 0]).float().contiguous()


def eligible(x, head_dim, heads, kv_heads, cfg, input_mask, alibi, layer_past):
    """True when the step can run on the arena kernels: bf16 on the GPU, head size 64 / 128, plain causal masking; ALiBi as
    per-head slopes (BLOOM), local windows (GPT-Neo).  Padding masks keep the general masked-softmax path."""
    if not (x.is_cuda and x.dtype == torch.bfloat16 and head_dim in (64, 128)) or os.environ.get("DSB200_V1_ARENA", "1") == "0":
        return False
    if heads % kv_heads or 128 % (heads // kv_heads):
        return False
    if layer_past is not None or not cfg.triangular_masking:
        return False
    if alibi is not None and (alibi.numel() < 2 * x.shape[0] * heads or x.shape[0] != 1 and input_mask is not None):
        return False  # slopes are read off the tensor: needs two key positions and unpadded sequences
    return input_mask is None or input_mask.numel() <= 1 or x.shape[0] == 1  # one sequence has no padding to mask


class KVArena:

    def __init__(self, min_tokens=1024):
        self.min_tokens = min_tokens
        self.cache = None
        self.block_table = None
        self.batch = self.per_seq = self.length = 0

    def reset(self):
        self.length = 0

    def _ensure(self, B, tokens, kv_heads, d, device):
        per_seq = -(-max(tokens, self.min_tokens) // BLOCK)
        fits = (self.cache is not None and self.batch == B and self.per_seq >= -(-tokens // BLOCK) and self.cache.shape[3] == kv_heads
                and self.cache.shape[4] == d and self.cache.device == device)
        if fits:
            return
        grow = self.cache is not None and self.batch == B and self.length > 0 and self.cache.shape[3:] == (kv_heads, d)
        if grow:
            per_seq = max(per_seq, 2 * self.per_seq)
        new = torch.zeros(B * per_seq, 2, BLOCK, kv_heads, d, dtype=torch.bfloat16, device=device)  # never-written slots must be finite
        if grow:
            new.view(B, per_seq, 2, BLOCK, kv_heads, d)[:, :self.per_seq].copy_(self.cache.view(B, self.per_seq, 2, BLOCK, kv_heads, d))
        else:
            self.length = 0
        self.cache, self.batch, self.per_seq = new, B, per_seq
        self.block_table = torch.arange(B * per_seq, device=device, dtype=torch.int32).view(B, per_seq)

    def attend(self, qkv, heads, kv_heads, d, scale, inv_freq=None, rot_dim=0, alibi_slopes=None, window=0):
        """qkv: [B, S, (heads + 2 kv_heads) * d], fresh projection output (rotated and consumed in place).
        Appends the S new rows of every sequence at positions [length, length + S) and returns the context [B, S, heads * d]."""
        from ....inference.v2 import kernels as K  # late: inference/ imports this package while it loads
        B, S, _ = qkv.shape
        self._ensure(B, self.length + S, kv_heads, d, qkv.device)
        flat = qkv.reshape(B * S, -1)
        if not flat.is_contiguous():
            flat = flat.contiguous()
        seq, pos = _token_index(B, S, self.length, qkv.device)
        if inv_freq is None:
            inv_freq, rot_dim = torch.empty(1, device=qkv.device, dtype=torch.float32), 0
        K.kv_rotary_append(flat, self.cache, seq, pos, self.block_table, inv_freq, heads, kv_heads, d, rot_dim)
        ctx = K.paged_flash_attention(flat, self.cache, seq, pos, self.block_table, heads, kv_heads, d, scale, max_context=self.length + S,
                                      alibi_slopes=alibi_slopes, window=window)
        self.length += S
        return ctx.view(B, S, heads * d)

    def keys_values(self):
        """The stored history as dense [B, kv_heads, T, d] tensors (for callers that ask for `present`)."""
        B, T = self.batch, self.length
        c = self.cache.view(B, self.per_seq, 2, BLOCK, *self.cache.shape[3:])
        k = c[:, :, 0].reshape(B, self.per_seq * BLOCK, *self.cache.shape[3:])[:, :T].transpose(1, 2)
        v = c[:, :, 1].reshape(B, self.per_seq * BLOCK, *self.cache.shape[3:])[:, :T].transpose(1, 2)
        return k, v
