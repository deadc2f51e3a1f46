"""ZeRO++ collectives: qwZ (quantized weight all-gather), qgZ (quantized gradient reduce-scatter), hpZ (secondary
shard group).  Parity: partition_parameters.py:679-717,1170-1266 (CUDAQuantizer + quantized gather),
runtime/comm/coalesced_collectives.py:31-68 (all_to_all_quant_reduce), utils/groups.py:505-532.

B200 design: weights travel as block-scaled fp8 (e4m3 payload + one fp32 scale per 128-element block) instead of the
reference's int8 -- same 2x volume reduction, and the payload is directly consumable by the fp8 tensor-core path.
Gradients use the same block-scaled fp8 format for the all-to-all based reduce-scatter.  On one NVSwitch domain the
reference's two-hop (intra-node, inter-node) qgZ collapses to one hop; the 2-level code path is kept for API parity.
"""
import torch

from ... import comm as dist
from ...ops import quantizer as Q


def quantized_all_gather(out_full, shard, group):
    """All-gather `shard` into `out_full` moving fp8 payload + scales (qwZ)."""
    world = dist.get_world_size(group)
    q, scales = Q.quantize_fp8_block(shard)
    q_all = torch.empty(q.numel() * world, dtype=q.dtype, device=q.device)
    s_all = torch.empty(scales.numel() * world, dtype=scales.dtype, device=scales.device)
    dist.all_gather_into_tensor(q_all, q, group=group)
    dist.all_gather_into_tensor(s_all, scales, group=group)
    Q.dequantize_fp8_block(q_all, s_all, out=out_full)
    return out_full


def quantized_reduce_scatter(out_shard, full_grad, group):
    """qgZ: reduce-scatter of `full_grad` ([world * n]) with fp8 block-scaled payload: quantize -> all-to-all ->
    dequant-reduce.  Result (sum over ranks of their slice r) lands in `out_shard` ([n])."""
    world = dist.get_world_size(group)
    n = full_grad.numel() // world
    q, scales = Q.quantize_fp8_block(full_grad)
    q_recv = torch.empty_like(q)
    s_recv = torch.empty_like(scales)
    dist.all_to_all_single(q_recv, q, group=group)
    dist.all_to_all_single(s_recv, scales, group=group)
    Q.dequant_reduce_fp8_block(q_recv, s_recv, world, out=out_shard)
    return out_shard


def all_to_all_quant_reduce(tensors, groups):
    """Reference-signature entry point (coalesced_collectives.py:31): list of gradient tensors -> list of this rank's
    reduced partitions."""
    group = groups.get("local_0") if isinstance(groups, dict) else groups
    world = dist.get_world_size(group)
    out = []
    for t in tensors:
        flat = t.reshape(-1)
        pad = (-flat.numel()) % (world * Q.FP8_BLOCK)
        if pad:
            flat = torch.cat([flat, flat.new_zeros(pad)])
        shard = torch.empty(flat.numel() // world, dtype=t.dtype, device=t.device)
        quantized_reduce_scatter(shard, flat, group)
        out.append(shard.div_(world))
    return out
