"""DeepSpeed-Ulysses sequence parallelism: all-to-all re-sharding (sequence <-> heads) around a local attention.

Parity: deepspeed/sequence/layer.py:15-109 (`single_all_to_all`, `_SeqAllToAll`, `DistributedAttention`).
Difference (the reference leaves it as a TODO at :96-101): q, k and v are exchanged in ONE all-to-all when they have
the same shape (or via `forward_packed` for a packed QKV projection), and the collective runs in
`seq_parallel_comm_data_type`.
"""
import torch

from .. import comm as dist


def single_all_to_all(input, scatter_idx, gather_idx, group):
    """Split dim `scatter_idx` into world-size chunks (chunk p goes to rank p) and concatenate the received chunks
    along dim `gather_idx` (rank-major).  Works for any tensor rank / batch size."""
    P = dist.get_world_size(group)
    if P == 1:
        return input
    assert input.shape[scatter_idx] % P == 0, f"dim {scatter_idx} ({input.shape[scatter_idx]}) not divisible by {P}"
    send = torch.stack(input.chunk(P, dim=scatter_idx), dim=0).contiguous()
    recv = torch.empty_like(send)
    dist.all_to_all_single(recv, send, group=group)
    out_shape = list(send.shape[1:])
    out_shape[gather_idx] *= P
    return recv.movedim(0, gather_idx).reshape(out_shape)


class _SeqAllToAll(torch.autograd.Function):

    @staticmethod
    def forward(ctx, group, input, scatter_idx, gather_idx):
        ctx.group = group
        ctx.scatter_idx = scatter_idx
        ctx.gather_idx = gather_idx
        return single_all_to_all(input, scatter_idx, gather_idx, group)

    @staticmethod
    def backward(ctx, *grad_output):
        return (None, _SeqAllToAll.apply(ctx.group, *grad_output, ctx.gather_idx, ctx.scatter_idx), None, None)


class DistributedAttention(torch.nn.Module):
    """Wraps any local attention `local_attn(q, k, v, *args)`.

    Inputs are sequence-sharded: [s/P, b, h, d] (scatter_idx=2, gather_idx=0 -- the reference default) or
    [b, s/P, h, d] (scatter_idx=2, gather_idx=1).  Inside, every rank holds the full sequence for h/P heads."""

    def __init__(self, local_attention, sequence_process_group, scatter_idx=2, gather_idx=0):
        super().__init__()
        self.local_attn = local_attention
        self.spg = sequence_process_group
        self.scatter_idx = scatter_idx
        self.gather_idx = gather_idx

    def forward(self, query, key, value, *args, **kwargs):
        other = [d for d in range(query.dim()) if d not in (self.scatter_idx, self.gather_idx)]
        if query.shape == key.shape == value.shape and other:
            # one fused exchange for q, k, v: concatenate along a dim that is neither scattered nor gathered
            cat_dim = other[0]
            qkv = torch.cat([query, key, value], dim=cat_dim)
            qkv = _SeqAllToAll.apply(self.spg, qkv, self.scatter_idx, self.gather_idx)
            query_layer, key_layer, value_layer = qkv.chunk(3, dim=cat_dim)
        else:  # GQA: k/v have fewer heads
            query_layer = _SeqAllToAll.apply(self.spg, query, self.scatter_idx, self.gather_idx)
            key_layer = _SeqAllToAll.apply(self.spg, key, self.scatter_idx, self.gather_idx)
            value_layer = _SeqAllToAll.apply(self.spg, value, self.scatter_idx, self.gather_idx)
        context_layer = self.local_attn(query_layer, key_layer, value_layer, *args, **kwargs)
        return _SeqAllToAll.apply(self.spg, context_layer, self.gather_idx, self.scatter_idx)
