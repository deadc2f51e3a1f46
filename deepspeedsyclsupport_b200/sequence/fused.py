"""DeepSpeed-Ulysses on a packed QKV projection, with the two all-to-alls done as one-sided peer-memory kernels that also do
the layout change (csrc/comm/symm.cu ulysses_heads_kernel):

    sequence-sharded  [B, S/P, (hq + 2 hkv) D]   --push-->   head-sharded  [B, S, (hq/P + 2 hkv/P) D]     (q, k, v in one pass)
    head-sharded      [B, S, (hq/P) D]           --pull-->   sequence-sharded  [B, S/P, hq D]             (attention output)

The reference does the same re-sharding with torch reshapes around `all_to_all_single` (sequence/layer.py:15-109, one call each
for q, k and v).  Here a row of the local tensor is stored straight into the owning ranks' attention input (no send-side
staging, no NCCL), and the attention output is read straight out of the owners' buffers.  Each exchange is the other's
backward.  Without peer memory (CPU / gloo, or multi-node groups) the same functions fall back to `single_all_to_all`.
"""
import ctypes
import os

import torch

from .. import comm as dist
from ..ops import native
from .layer import single_all_to_all

native.register_sigs({"dsb200_ulysses_heads_exchange": "p l p iiii iii p ii i p"})


class _Exchange:
    """Symmetric buffers of one sequence-parallel group: region 0 receives pushes / serves pulls of the forward pass, region 1
    does the same for gradients (forward and backward of neighbouring layers may be in flight on different ranks)."""
    _by_group = {}

    def __init__(self, group, nbytes):
        from ..parallel.symm import SymmetricBuffer
        self.group = group
        self.region = (int(nbytes) + 255) // 256 * 256
        self.symm = SymmetricBuffer(group, 2 * self.region)
        self.world, self.rank = self.symm.world, self.symm.rank

    @classmethod
    def get(cls, group, nbytes):
        group = group if group is not None else dist.get_world_group()
        key = tuple(dist.get_all_ranks_from_group(group))
        cur = cls._by_group.get(key)
        if cur is None or cur.region < nbytes:
            if cur is not None:
                cur.symm.close()
            cur = cls._by_group[key] = cls(group, nbytes)
        return cur

    def run(self, region, local, B, S_local, D, src_ld, dst_ld, segs, push):
        flat = (ctypes.c_int * (4 * len(segs)))(*[v for seg in segs for v in seg])
        tab = ctypes.cast((ctypes.c_void_p * self.world)(*self.symm.ptrs), ctypes.c_void_p)
        rc = native.lib().dsb200_ulysses_heads_exchange(tab, region * self.region, native.ptr(local), B, S_local, self.rank, self.world, D,
                                                        src_ld, dst_ld, ctypes.cast(flat, ctypes.c_void_p), len(segs), int(push), 0,
                                                        native.stream_ptr())
        native.check(rc)

    def view(self, region, rows, cols):
        return self.symm.view(region * self.region, (rows, cols), torch.bfloat16)


def peer_usable(x, group):
    P = dist.get_world_size(group) if dist.is_initialized() else 1
    return (P > 1 and x.is_cuda and x.dtype == torch.bfloat16 and os.environ.get("DSB200_ULYSSES_PEER", "1") != "0" and
            os.environ.get("DSB200_DISABLE_P2P", "0") != "1" and P <= torch.cuda.device_count())


def _segments(heads, D, P):
    """[(src_col, dst_col, heads, heads_per_rank)] for packed segments with `heads` = (hq, hkv, hkv) or (hq, )."""
    segs, src, dst = [], 0, 0
    for h in heads:
        assert h % P == 0, f"{h} heads do not split over {P} sequence-parallel ranks"
        segs.append((src, dst, h, h // P))
        src += h * D
        dst += h // P * D
    return segs, src, dst


class _ToHeads(torch.autograd.Function):
    """x [B * S_local, sum(heads) * D] (sequence-sharded) -> [B * S, sum(heads) / P * D] (head-sharded, whole sequence)."""

    @staticmethod
    def forward(ctx, x, B, S_local, heads, D, group):
        P = dist.get_world_size(group)
        segs, src_ld, dst_ld = _segments(heads, D, P)
        x = x.contiguous()
        rows = B * S_local * P
        ex = _Exchange.get(group, rows * dst_ld * 2)
        ex.symm.barrier()                      # everyone is done reading region 0 from the previous exchange
        ex.run(0, x, B, S_local, D, src_ld, dst_ld, segs, push=True)
        ex.symm.barrier()                      # every row has landed
        ctx.meta = (B, S_local, heads, D, group, src_ld, dst_ld, segs)
        return ex.view(0, rows, dst_ld).clone()

    @staticmethod
    def backward(ctx, g):
        B, S_local, heads, D, group, src_ld, dst_ld, segs = ctx.meta
        P = dist.get_world_size(group)
        rows = B * S_local * P
        ex = _Exchange.get(group, rows * dst_ld * 2)
        ex.symm.barrier()
        ex.view(1, rows, dst_ld).copy_(g.reshape(rows, dst_ld))
        ex.symm.barrier()
        gx = torch.empty(B * S_local, src_ld, dtype=torch.bfloat16, device=g.device)
        ex.run(1, gx, B, S_local, D, src_ld, dst_ld, segs, push=False)
        return gx, None, None, None, None, None


class _ToSequence(torch.autograd.Function):
    """y [B * S, heads / P * D] (head-sharded) -> [B * S_local, heads * D] (sequence-sharded)."""

    @staticmethod
    def forward(ctx, y, B, S_local, heads, D, group):
        P = dist.get_world_size(group)
        segs, src_ld, dst_ld = _segments((heads, ), D, P)
        rows = B * S_local * P
        ex = _Exchange.get(group, rows * dst_ld * 2)
        ex.symm.barrier()
        ex.view(0, rows, dst_ld).copy_(y.reshape(rows, dst_ld))
        ex.symm.barrier()
        out = torch.empty(B * S_local, src_ld, dtype=torch.bfloat16, device=y.device)
        ex.run(0, out, B, S_local, D, src_ld, dst_ld, segs, push=False)
        ctx.meta = (B, S_local, D, group, src_ld, dst_ld, segs)
        return out

    @staticmethod
    def backward(ctx, g):
        B, S_local, D, group, src_ld, dst_ld, segs = ctx.meta
        P = dist.get_world_size(group)
        rows = B * S_local * P
        ex = _Exchange.get(group, rows * dst_ld * 2)
        g = g.contiguous()
        ex.symm.barrier()
        ex.run(1, g, B, S_local, D, src_ld, dst_ld, segs, push=True)
        ex.symm.barrier()
        return ex.view(1, rows, dst_ld).clone(), None, None, None, None, None


def qkv_to_heads(qkv, B, S_local, hq, hkv, D, group):
    """Packed QKV of this rank's tokens -> packed QKV of this rank's heads over the whole sequence."""
    if peer_usable(qkv, group):
        return _ToHeads.apply(qkv, B, S_local, (hq, hkv, hkv), D, group)
    P = dist.get_world_size(group)
    parts = qkv.view(B, S_local, hq + 2 * hkv, D).split([hq, hkv, hkv], dim=2)
    out = [single_all_to_all(t.contiguous(), 2, 1, group) for t in parts]  # [B, S, h/P, D] each
    return torch.cat(out, dim=2).reshape(B * S_local * P, -1)


def attn_to_sequence(y, B, S_local, hq, D, group):
    """Attention output of this rank's heads over the whole sequence -> all heads for this rank's tokens."""
    if peer_usable(y, group):
        return _ToSequence.apply(y, B, S_local, hq, D, group)
    P = dist.get_world_size(group)
    out = single_all_to_all(y.view(B, S_local * P, hq // P, D).contiguous(), 1, 2, group)  # [B, S_local, hq, D]
    return out.reshape(B * S_local, hq * D)
