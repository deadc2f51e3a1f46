"""Batched collectives over lists of tensors (parity: deepspeed/runtime/comm/coalesced_collectives.py:29-139)."""
import math
from typing import List

import torch
from torch import Tensor

from ... import comm as dist


def reduce_scatter_coalesced(tensors: List[Tensor], group=None) -> List[Tensor]:
    """Simultaneously reduce-scatter a list of tensors - this can be done more efficiently than individual reduce scatter
    calls: every tensor is padded to a multiple of world, the rank-r slices are packed into one flat send buffer per rank
    and ONE reduce_scatter moves everything.  Returns this rank's (averaged) partition of each tensor."""
    this_rank = dist.get_rank(group)
    world_sz = dist.get_world_size(group)
    partition_lst_for_each_tensor = [None] * len(tensors)
    for tensor_idx, tensor in enumerate(tensors):
        flattened_tensor = tensor.view(-1)
        chunk_sz = math.ceil(tensor.numel() / world_sz)
        partition_lst_for_each_tensor[tensor_idx] = [flattened_tensor[rank * chunk_sz:rank * chunk_sz + chunk_sz] for rank in range(0, world_sz)]
    padded_partition_sz_for_each_tensor = tuple(math.ceil(t.numel() / world_sz) for t in tensors)
    if len(tensors) == 1 and tensors[0].numel() % world_sz == 0:
        tensor_partition_flat_buffer = tensors[0].view(-1)
    else:
        tensor_partitions_lst_with_padding = []
        for rank in range(world_sz):
            for tensor_idx in range(len(tensors)):
                tensor_chunk = partition_lst_for_each_tensor[tensor_idx][rank]
                tensor_partitions_lst_with_padding.append(tensor_chunk)
                padding_sz = padded_partition_sz_for_each_tensor[tensor_idx] - tensor_chunk.numel()
                if padding_sz > 0:
                    tensor_partitions_lst_with_padding.append(torch.empty(padding_sz, dtype=tensor_chunk.dtype, device=tensor_chunk.device))
        tensor_partition_flat_buffer = torch.cat(tensor_partitions_lst_with_padding)
    tensor_partition_flat_buffer.div_(world_sz)
    per_rank = tensor_partition_flat_buffer.numel() // world_sz
    out = torch.empty(per_rank, dtype=tensor_partition_flat_buffer.dtype, device=tensor_partition_flat_buffer.device)
    dist.reduce_scatter_tensor(out, tensor_partition_flat_buffer, group=group)
    output_lst = [None] * len(tensors)
    offset = 0
    for tensor_idx in range(len(tensors)):
        output_lst[tensor_idx] = out.narrow(0, offset, partition_lst_for_each_tensor[tensor_idx][this_rank].numel())
        offset += padded_partition_sz_for_each_tensor[tensor_idx]
    return output_lst


def all_to_all_quant_reduce(tensors: List[Tensor], groups: {}) -> List[Tensor]:
    """ZeRO++ qgZ quantized gradient reduce (parity: coalesced_collectives.py:29-73); implemented in runtime/zero/zeropp.py
    on the fp8-block kernels."""
    from ..zero.zeropp import all_to_all_quant_reduce as _impl
    return _impl(tensors, groups)
