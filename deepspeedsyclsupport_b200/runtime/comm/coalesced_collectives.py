"""Batched collectives over lists of tensors (parity: deepspeed/runtime/comm/coalesced_collectives.py:29-139)."""
from typing import List

import torch
from torch import Tensor

from ... import comm as dist


def reduce_scatter_coalesced(tensors: List[Tensor], group=None) -> List[Tensor]:
    """One reduce-scatter for a list of tensors; returns this rank's (averaged) partition of each.

    Tensor i is cut into `world` chunks of c_i = ceil(n_i / world) elements (the last ranks' chunks may be short or empty).  The
    send buffer is a [world, sum_i c_i] matrix: row r holds chunk r of every tensor, zero-padded to c_i, so ONE reduce_scatter
    of its rows delivers every partition.  A single tensor that divides evenly is sent (and averaged) in place."""
    rank, world = dist.get_rank(group), dist.get_world_size(group)
    chunk = [-(-t.numel() // world) for t in tensors]
    if len(tensors) == 1 and tensors[0].numel() % world == 0:
        send = tensors[0].view(-1)
    else:
        send = tensors[0].new_zeros(world, sum(chunk))
        col = 0
        for t, c in zip(tensors, chunk):
            flat = t.reshape(-1)
            whole, rest = divmod(flat.numel(), c) if c else (0, 0)
            send[:whole, col:col + c] = flat[:whole * c].view(whole, c)
            if rest:
                send[whole, col:col + rest] = flat[whole * c:]
            col += c
        send = send.view(-1)
    send.div_(world)
    mine = send.new_empty(send.numel() // world)
    dist.reduce_scatter_tensor(mine, send, group=group)
    parts, col = [], 0
    for t, c in zip(tensors, chunk):
        parts.append(mine.narrow(0, col, max(0, min(c, t.numel() - rank * c))))
        col += c
    return parts


def all_to_all_quant_reduce(tensors: List[Tensor], groups: {}) -> List[Tensor]:
    """ZeRO++ qgZ quantized gradient reduce (parity: coalesced_collectives.py:29-73); implemented in runtime/zero/zeropp.py
    on the fp8-block kernels."""
    from ..zero.zeropp import all_to_all_quant_reduce as _impl
    return _impl(tensors, groups)
