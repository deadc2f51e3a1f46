"""Python wrappers of the ragged device ops (csrc/ragged/ragged_ops.cu) with PyTorch oracles for CPU / tests."""
import math
import os

import torch

from ....ops import native

native.register_sigs({
    "dsb200_kv_rotary_append": "pppppp iiiiiii p",
    "dsb200_paged_attention": "pppppp iiiiii f p",
    "dsb200_paged_flash_attention": "pppppp iiiiiii f ii p p i p",
    "dsb200_paged_flash_decode": "pppppp iiiiiii f ii p p i p",
    "dsb200_paged_flash_scratch_bytes": "iiiii",
    "dsb200_ragged_embed": "ppppp iii p",
    "dsb200_logits_gather": "ppp ii p",
    "dsb200_top_k_gating": "ppppp iiii p",
    "dsb200_moe_scatter": "ppppppp iiii p",
    "dsb200_moe_gather": "pppp iii p",
}, restype_long=("dsb200_paged_flash_scratch_bytes", ))


def kv_rotary_append(qkv, kv_cache, tok_seq, tok_pos, block_table, inv_freq, hq, hkv, d, rot_dim=None):
    """In place: RoPE on q,k heads of qkv [T, (hq+2hkv)*d]; k,v rows appended to kv_cache [blocks,2,bs,hkv,d]."""
    T = qkv.shape[0]
    bs = kv_cache.shape[2]
    rot_dim = d if rot_dim is None else rot_dim
    if qkv.is_cuda:
        rc = native.lib().dsb200_kv_rotary_append(native.ptr(qkv), native.ptr(kv_cache), native.ptr(tok_seq), native.ptr(tok_pos),
                                                  native.ptr(block_table), native.ptr(inv_freq), T, hq, hkv, d, bs,
                                                  block_table.shape[1], int(rot_dim), native.stream_ptr())
        native.check(rc)
        return qkv
    from ....ops.fused_ops import rope_ref
    if rot_dim == d:
        qkv.copy_(rope_ref(qkv, inv_freq, tok_pos, 0, hq + hkv, d))
    elif rot_dim > 0:
        xv = qkv.view(T, hq + 2 * hkv, d)
        rot = rope_ref(xv[:, :hq + hkv, :rot_dim].reshape(T, -1).contiguous(), inv_freq, tok_pos, 0, hq + hkv, rot_dim)
        xv[:, :hq + hkv, :rot_dim] = rot.view(T, hq + hkv, rot_dim)
    x = qkv.view(T, hq + 2 * hkv, d)
    for t in range(T):
        pos, seq = int(tok_pos[t]), int(tok_seq[t])
        blk = int(block_table[seq, pos // bs])
        kv_cache[blk, 0, pos % bs] = x[t, hq:hq + hkv]
        kv_cache[blk, 1, pos % bs] = x[t, hq + hkv:]
    return qkv


def paged_attention_ref(qkv, kv_cache, tok_seq, tok_pos, block_table, hq, hkv, d, scale, alibi_slopes=None, window=0):
    T = qkv.shape[0]
    bs = kv_cache.shape[2]
    out = torch.empty(T, hq * d, dtype=qkv.dtype, device=qkv.device)
    q = qkv.view(T, hq + 2 * hkv, d)[:, :hq].float()
    for t in range(T):
        n = int(tok_pos[t]) + 1
        seq = int(tok_seq[t])
        blocks = block_table[seq, :-(-n // bs)].long()
        k = kv_cache[blocks, 0].reshape(-1, hkv, d)[:n].float().repeat_interleave(hq // hkv, 1)
        v = kv_cache[blocks, 1].reshape(-1, hkv, d)[:n].float().repeat_interleave(hq // hkv, 1)
        s = torch.einsum("hd,nhd->hn", q[t], k) * scale
        if alibi_slopes is not None:
            s = s + alibi_slopes.float().to(s.device)[:, None] * torch.arange(n, device=s.device, dtype=torch.float32)[None, :]
        if window and n > window:
            s[:, :n - window] = float("-inf")
        out[t] = torch.einsum("hn,nhd->hd", torch.softmax(s, -1), v).reshape(-1).to(qkv.dtype)
    return out


def paged_flash_supported(qkv, kv_cache, hq, hkv, d):
    """The tcgen05 kernel covers bf16, head_dim 64 / 128, q:kv head ratios and KV block sizes that divide 128."""
    group = hq // max(hkv, 1)
    return (qkv.is_cuda and qkv.dtype == torch.bfloat16 and kv_cache.dtype == torch.bfloat16 and d in (64, 128) and hq % hkv == 0 and
            128 % group == 0 and kv_cache.shape[2] >= 8 and 128 % kv_cache.shape[2] == 0 and os.environ.get("DSB200_PAGED_FLASH", "1") != "0")


def _pick_splits(T, hq, d, ctas, pages, page_us=0.7):
    """KV splits per (atom, kv head): more splits fill the 148 SMs when the batch is small, but every split writes an fp32
    partial per (token, head) that the merge kernel reads back.  Cost model in microseconds: waves of CTAs x (5 us of set-up + pages per CTA at
    ~0.7 us a page), plus the partial traffic at HBM speed and the merge launch."""
    best, best_cost = 1, None
    for n in range(1, min(16, pages) + 1):
        waves = -(-ctas * n // 148)
        cost = (5.0 + -(-pages // n) * page_us) * waves
        if n > 1:
            cost += 2.0 * T * hq * (d + 2) * 4 * n / 6.0e6 + 3.0
        if best_cost is None or cost < best_cost - 1e-9:
            best, best_cost = n, cost
    return best


def paged_flash_attention(qkv, kv_cache, tok_seq, tok_pos, block_table, hq, hkv, d, scale=None, max_context=None, n_splits=None,
                          decode=None, alibi_slopes=None, window=0):
    """Attention of every ragged token over its sequence's paged KV history on the tensor cores (csrc/attn/paged_flash.cu):
    atoms of up to 128/group tokens x kv head, KV pages fetched by TMA through the block table, prefill chunks and decode steps
    alike; small batches over long contexts are split along the KV axis (`n_splits`, chosen here unless given) and merged.
    `decode`: the batch is (almost) one token per sequence -- use the transposed small-atom kernel (keys on the MMA M axis,
    16 q rows per atom).  Default: when the batch has no more tokens than the block table has sequences and d == 128.
    `alibi_slopes`: optional [hq] fp32 device tensor; adds slope[head] * key position to the scores (BLOOM-style ALiBi).
    `window` > 0: local attention over the last `window` keys (GPT-Neo's local layers)."""
    scale = scale if scale is not None else 1.0 / math.sqrt(d)
    T = qkv.shape[0]
    bs, group = kv_cache.shape[2], hq // hkv
    if decode is None:
        decode = T <= block_table.shape[0]
    decode = bool(decode) and d == 128 and 16 % group == 0 and os.environ.get("DSB200_PAGED_DECODE", "1") != "0"
    nt = (16 if decode else 128) // group
    max_atoms = max(1, min(T, (T + nt - 1) // nt + block_table.shape[0]))
    if n_splits is None:
        pages = -(-(max_context if max_context is not None else block_table.shape[1] * bs) // 128)
        n_splits = _pick_splits(T, hq, d, max_atoms * hkv, pages, page_us=0.45 if decode else 0.7)
    out = torch.empty(T, hq * d, dtype=qkv.dtype, device=qkv.device)
    nbytes = native.lib().dsb200_paged_flash_scratch_bytes(T, hq, d, max_atoms, n_splits)
    scratch = torch.empty(nbytes, dtype=torch.uint8, device=qkv.device)
    fn = native.lib().dsb200_paged_flash_decode if decode else native.lib().dsb200_paged_flash_attention
    rc = fn(native.ptr(qkv), native.ptr(kv_cache), native.ptr(out), native.ptr(tok_seq), native.ptr(tok_pos), native.ptr(block_table), T, hq,
            hkv, d, bs, block_table.shape[1], kv_cache.shape[0], float(scale), max_atoms, n_splits, native.ptr(scratch),
            native.ptr(alibi_slopes), int(window), native.stream_ptr())
    native.check(rc)
    return out


def paged_attention(qkv, kv_cache, tok_seq, tok_pos, block_table, hq, hkv, d, scale=None, max_context=None, decode=None, alibi_slopes=None,
                    window=0):
    scale = scale if scale is not None else 1.0 / math.sqrt(d)
    if not qkv.is_cuda:
        return paged_attention_ref(qkv, kv_cache, tok_seq, tok_pos, block_table, hq, hkv, d, scale, alibi_slopes, window)
    if paged_flash_supported(qkv, kv_cache, hq, hkv, d):
        return paged_flash_attention(qkv, kv_cache, tok_seq, tok_pos, block_table, hq, hkv, d, scale, max_context=max_context, decode=decode,
                                     alibi_slopes=alibi_slopes, window=window)
    assert alibi_slopes is None and not window, "ALiBi / local attention need the tensor-core kernel (bf16, head size 64 / 128)"
    # head sizes the tensor-core kernel does not cover (96, 256): CUDA-core kernel, one CTA per (token, kv head)
    T = qkv.shape[0]
    out = torch.empty(T, hq * d, dtype=qkv.dtype, device=qkv.device)
    rc = native.lib().dsb200_paged_attention(native.ptr(qkv), native.ptr(kv_cache), native.ptr(out), native.ptr(tok_seq),
                                             native.ptr(tok_pos), native.ptr(block_table), T, hq, hkv, d, kv_cache.shape[2],
                                             block_table.shape[1], float(scale), native.stream_ptr())
    native.check(rc)
    return out


def ragged_embed(ids, wte, wpe=None, tok_pos=None, pos_offset=0):
    if not wte.is_cuda or wte.dtype != torch.bfloat16:
        x = wte[ids]
        if wpe is not None:
            x = x + wpe[tok_pos.long() + pos_offset]
        return x
    T, dim = ids.numel(), wte.shape[1]
    out = torch.empty(T, dim, dtype=wte.dtype, device=wte.device)
    rc = native.lib().dsb200_ragged_embed(native.ptr(ids), native.ptr(wte), native.ptr(wpe), native.ptr(tok_pos if wpe is not None else None),
                                          native.ptr(out), T, dim, pos_offset, native.stream_ptr())
    native.check(rc)
    return out


def logits_gather(hidden, last_tok):
    if not hidden.is_cuda or hidden.dtype != torch.bfloat16:
        return hidden[last_tok.long()]
    out = torch.empty(last_tok.numel(), hidden.shape[1], dtype=hidden.dtype, device=hidden.device)
    rc = native.lib().dsb200_logits_gather(native.ptr(hidden), native.ptr(last_tok), native.ptr(out), last_tok.numel(), hidden.shape[1],
                                           native.stream_ptr())
    native.check(rc)
    return out


def top_k_gating(logits, k, normalize=True):
    """logits [T, E] fp32 -> (expert_counts [E] int32, scores [T,k] fp32, assignments [T,k] int32, offsets [T,k] int32)."""
    T, E = logits.shape
    if not logits.is_cuda:
        probs = torch.softmax(logits.float(), -1)
        sc, asg = probs.topk(k, dim=-1)
        if normalize:
            sc = sc / sc.sum(-1, keepdim=True)
        counts = torch.zeros(E, dtype=torch.int32)
        offsets = torch.zeros(T, k, dtype=torch.int32)
        for t in range(T):
            for j in range(k):
                e = int(asg[t, j])
                offsets[t, j] = counts[e]
                counts[e] += 1
        return counts, sc, asg.int(), offsets
    counts = torch.zeros(E, dtype=torch.int32, device=logits.device)
    scores = torch.empty(T, k, dtype=torch.float32, device=logits.device)
    asg = torch.empty(T, k, dtype=torch.int32, device=logits.device)
    off = torch.empty(T, k, dtype=torch.int32, device=logits.device)
    rc = native.lib().dsb200_top_k_gating(native.ptr(logits.contiguous()), native.ptr(counts), native.ptr(scores), native.ptr(asg),
                                          native.ptr(off), T, E, k, int(normalize), native.stream_ptr())
    native.check(rc)
    return counts, scores, asg, off


def moe_scatter(x, counts, assignments, offsets):
    """rows of x copied into expert-sorted order.  returns (moe_in [T*k, dim], expert_cumsum [E] int64, mapped_slots [T,k])."""
    T, dim = x.shape
    k = assignments.shape[1]
    E = counts.numel()
    if not x.is_cuda or x.dtype != torch.bfloat16:
        cumsum = counts.long().cumsum(0)
        start = torch.cat([cumsum.new_zeros(1), cumsum[:-1]])
        slots = (start[assignments.long()] + offsets.long()).int()
        moe_in = x.new_zeros(T * k, dim)
        for j in range(k):
            moe_in[slots[:, j].long()] = x
        return moe_in, cumsum, slots
    moe_in = torch.empty(T * k, dim, dtype=x.dtype, device=x.device)
    cumsum = torch.empty(E, dtype=torch.int64, device=x.device)
    slots = torch.empty(T, k, dtype=torch.int32, device=x.device)
    rc = native.lib().dsb200_moe_scatter(native.ptr(x), native.ptr(counts), native.ptr(assignments), native.ptr(offsets),
                                         native.ptr(cumsum), native.ptr(moe_in), native.ptr(slots), T, E, dim, k, native.stream_ptr())
    native.check(rc, 2)
    return moe_in, cumsum, slots


def moe_gather(moe_out, scores, mapped_slots):
    T, k = mapped_slots.shape
    dim = moe_out.shape[1]
    if not moe_out.is_cuda or moe_out.dtype != torch.bfloat16:
        return (moe_out[mapped_slots.long()].float() * scores[..., None]).sum(1).to(moe_out.dtype)
    out = torch.empty(T, dim, dtype=moe_out.dtype, device=moe_out.device)
    rc = native.lib().dsb200_moe_gather(native.ptr(moe_out), native.ptr(scores), native.ptr(mapped_slots), native.ptr(out), T, dim, k,
                                        native.stream_ptr())
    native.check(rc)
    return out
