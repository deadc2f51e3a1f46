"""Fused compute + collective ops built on peer memory.

* `ag_gemm`   : Y = X W^T with the all-gather of the sharded W performed INSIDE the tcgen05 GEMM kernel
                (csrc/gemm/gemm_ag.cu).
* `rs_adam`   : reduce-scatter of a gradient buffer fused with the partitioned AdamW update (parallel/symm.py).
* `gemm_allreduce`: row-parallel GEMM followed by a one-shot peer-memory all-reduce (tensor-parallel inference).
"""
import ctypes

import torch

from ..ops import native
from ..ops.gemm import gemm

native.register_sigs({
    "dsb200_gemm_ag_bf16": "ppp iiiii p ll p i p",
})


def ag_gemm(x, w_local, out, peer_shard_ptrs, w_off, shard_numel, flags, N, K):
    """x: [M,K] bf16; w_local: [N,K] bf16 buffer that receives the gathered weight; peer_shard_ptrs: list of device
    pointers to every rank's shard region (own included); flags: int32 scratch with >= ceil(N/256) entries."""
    M = x.shape[0]
    world = len(peer_shard_ptrs)
    arr = (ctypes.c_void_p * world)(*peer_shard_ptrs)
    rc = native.lib().dsb200_gemm_ag_bf16(native.ptr(x), native.ptr(w_local), native.ptr(out), M, N, K, x.stride(0), out.stride(0),
                                          ctypes.cast(arr, ctypes.c_void_p), w_off, shard_numel, native.ptr(flags), world,
                                          native.stream_ptr())
    native.check(rc)
    return out


def gemm_allreduce(x, w_shard, symm, offset=0, out=None):
    """Row-parallel linear for tensor-parallel inference: partial = x @ w_shard^T written into the symmetric buffer,
    device barrier, one-shot all-reduce by peer loads.  Returns the reduced [M, N] bf16 tensor."""
    M, N = x.shape[0], w_shard.shape[0]
    partial = symm.view(offset, (M, N), torch.bfloat16)
    gemm(x, w_shard, out=partial)
    symm.barrier()
    if out is None:
        out = torch.empty(M, N, dtype=torch.bfloat16, device=x.device)
    symm.all_reduce_bf16(out, offset, M * N)
    symm.barrier()  # nobody overwrites its partial before every peer has read it
    return out
