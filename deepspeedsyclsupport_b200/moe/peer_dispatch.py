"""Expert-parallel dispatch / combine over NVLink peer memory (one NVSwitch domain) for MoE TRAINING.

The reference scatters tokens into an [E, C, M] buffer, all-to-alls it, runs the local experts, all-to-alls the outputs back
and gathers them (moe/sharded_moe.py:477-554: einsum + `_AllToAll` twice).  Here both exchanges are one-sided kernels
(csrc/comm/symm.cu):

    dispatch   every (token, k) row is STORED straight into the owning rank's receive buffer [src rank][local expert][slot]
               -- scatter and all-to-all in one pass over the tokens, no intermediate [E, C, M] tensor on the sender
    combine    every token PULLS its k expert-output rows from the owners and sums them with the gate weights
               -- all-to-all and gather in one pass, nothing but the needed rows crosses NVLink

Both are autograd Functions: the backward of dispatch is an unweighted pull of the gradient rows, the backward of combine is a
weight-scaled push plus the gate-weight gradient <expert row, grad row> computed while pulling.  Cross-rank ordering uses the
symmetric buffer's device barrier (no host synchronisation, no NCCL).  Two symmetric regions of [W][n_local][C][M] bf16 are
reused by every MoE layer of the model; data that must outlive the next layer (expert inputs saved for wgrad) is cloned out.
"""
import ctypes
import os

import torch

from .. import comm as dist
from ..ops import native

native.register_sigs({
    "dsb200_moe_dispatch_push": "p l pppp iii ii ii p",
    "dsb200_moe_combine_pull": "p l pppppp iii ii ii p",
})


class PeerMoEBuffers:
    """Symmetric regions A (rows arriving at this rank) and B (rows this rank serves to pullers) for one expert-parallel group."""
    _by_group = {}

    def __init__(self, group, n_local, capacity, model_dim):
        from ..parallel.symm import SymmetricBuffer
        self.group, self.n_local, self.capacity, self.M = group, n_local, capacity, model_dim
        self.world, self.rank = dist.get_world_size(group), dist.get_rank(group)
        self.rows = self.world * n_local * capacity
        self.region_bytes = (self.rows * model_dim * 2 + 255) // 256 * 256
        self.symm = SymmetricBuffer(group, 2 * self.region_bytes)
        self.off = (0, self.region_bytes)
        self.views = [self.symm.view(o, (self.rows, model_dim), torch.bfloat16) for o in self.off]

    @classmethod
    def get(cls, group, n_local, capacity, model_dim):
        group = group if group is not None else dist.get_world_group()
        key = tuple(dist.get_all_ranks_from_group(group))
        cur = cls._by_group.get(key)
        if cur is None or cur.n_local != n_local or cur.M != model_dim or cur.capacity < capacity:
            if cur is not None:
                cur.symm.close()
            cur = cls._by_group[key] = cls(group, n_local, capacity, model_dim)  # collective: every rank sees the same sizes
        return cur

    def _table(self):
        return ctypes.cast((ctypes.c_void_p * self.world)(*self.symm.ptrs), ctypes.c_void_p)

    def push(self, region, x, token_expert, token_slot, row_scale=None):
        S, K = token_expert.shape
        rc = native.lib().dsb200_moe_dispatch_push(self._table(), self.off[region], native.ptr(x), native.ptr(token_expert),
                                                   native.ptr(token_slot), native.ptr(row_scale), S, K, self.M, self.rank, self.world,
                                                   self.n_local, self.capacity, native.stream_ptr())
        native.check(rc)

    def pull(self, region, out, token_expert, token_slot, weights=None, dot_with=None, dot_out=None):
        S, K = token_expert.shape
        rc = native.lib().dsb200_moe_combine_pull(self._table(), self.off[region], native.ptr(out), native.ptr(token_expert),
                                                  native.ptr(token_slot), native.ptr(weights), native.ptr(dot_with), native.ptr(dot_out),
                                                  S, K, self.M, self.rank, self.world, self.n_local, self.capacity, native.stream_ptr())
        native.check(rc)


def usable(x, ep_size, group):
    return (ep_size > 1 and x.is_cuda and x.dtype == torch.bfloat16 and x.shape[-1] % 8 == 0 and dist.is_initialized() and
            dist.get_world_size(group) == ep_size and os.environ.get("DSB200_MOE_PEER", "1") != "0" and os.environ.get("DSB200_DISABLE_P2P", "0") != "1" and
            ep_size <= torch.cuda.device_count())


def _idx(t):
    return t.to(torch.int32).contiguous()


class PeerDispatch(torch.autograd.Function):
    """x [S, M] -> the rows this rank's experts received, [W, n_local, C, M] (zero where no token was routed)."""

    @staticmethod
    def forward(ctx, x, token_expert, token_slot, bufs):
        te, ts = _idx(token_expert), _idx(token_slot)
        x = x.contiguous()
        A = bufs.views[0]
        A.zero_()
        bufs.symm.barrier()            # every receive buffer is clear before anyone stores into it
        bufs.push(0, x, te, ts)
        bufs.symm.barrier()            # all rows have landed
        ctx.bufs, ctx.S = bufs, x.shape[0]
        ctx.save_for_backward(te, ts)
        return A.clone().view(bufs.world, bufs.n_local, bufs.capacity, bufs.M)

    @staticmethod
    def backward(ctx, grad_recv):
        te, ts = ctx.saved_tensors
        bufs = ctx.bufs
        B = bufs.views[1]
        B.copy_(grad_recv.reshape(bufs.rows, bufs.M))
        bufs.symm.barrier()            # every rank's gradient rows are staged
        gx = torch.empty(ctx.S, bufs.M, dtype=torch.bfloat16, device=grad_recv.device)
        bufs.pull(1, gx, te, ts)       # dropped tokens (expert < 0) contribute nothing: the kernel writes their sum of zero rows
        bufs.symm.barrier()            # staging region free again
        return gx, None, None, None


class PeerCombine(torch.autograd.Function):
    """expert_out [W, n_local, C, M] (this rank's experts' rows) -> y [S, M] = sum_k w[s, k] * row(expert_k, slot_k)."""

    @staticmethod
    def forward(ctx, expert_out, token_expert, token_slot, token_weight, S, bufs):
        te, ts = _idx(token_expert), _idx(token_slot)
        w = token_weight.float().contiguous()
        B = bufs.views[1]
        B.copy_(expert_out.reshape(bufs.rows, bufs.M))
        bufs.symm.barrier()
        y = torch.empty(S, bufs.M, dtype=torch.bfloat16, device=expert_out.device)
        bufs.pull(1, y, te, ts, weights=w)
        bufs.symm.barrier()
        ctx.bufs = bufs
        ctx.save_for_backward(expert_out, te, ts, w)
        return y

    @staticmethod
    def backward(ctx, gy):
        expert_out, te, ts, w = ctx.saved_tensors
        bufs = ctx.bufs
        gy = gy.contiguous()
        A, B = bufs.views
        A.zero_()
        B.copy_(expert_out.reshape(bufs.rows, bufs.M))
        bufs.symm.barrier()
        bufs.push(0, gy, te, ts, row_scale=w)                       # d expert_out rows = w * gy[s], stored at their owners
        gw = torch.zeros_like(w)
        bufs.pull(1, None, te, ts, dot_with=gy, dot_out=gw)         # d w[s, k] = <expert row, gy[s]>
        bufs.symm.barrier()
        g_expert = A.clone().view_as(expert_out)
        return g_expert, None, None, gw, None, None
