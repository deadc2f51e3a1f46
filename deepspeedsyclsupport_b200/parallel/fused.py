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
    "dsb200_gemm_ar_bf16": "ppp ll i pp iii iiii ii i p",
    "dsb200_gemm_ar_bf16_dev": "ppp lll p pp iii iiii ii ii i p",
    "dsb200_gemm_rs_bf16": "ppp ll i pp iii iiii ii ii i p",
})


def ag_gemm(x, w_local, out, peer_shard_ptrs, w_off, shard_numel, flags, N, K):
    """x: [M,K] bf16; w_local: [N,K] bf16 buffer that receives the gathered weight; peer_shard_ptrs: list of device
    pointers to every rank's shard region (own included); flags: int32 scratch with >= ceil(N/256) + 1 entries."""
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


class FusedGemmAllReduce(object):
    """Row-parallel linear + all-reduce as ONE kernel per call (csrc/gemm/gemm_ar.cu): every rank's GEMM epilogue stores its
    partial tiles into its own symmetric buffer and raises a per-tile flag; reducer warps of the same kernel pull the peers'
    tiles as their flags arrive and write the sum, tile by tile under the remaining MMAs.

    `symm`: a SymmetricBuffer (or EmulatedPeers with world 1) holding, from `offset`: two partial areas of max_m * n * 2 bytes
    (double-buffered by call parity) followed by the per-tile flags.  Every rank must make the same sequence of calls."""

    def __init__(self, symm, max_m, n, offset=0, device_epochs=False):
        """device_epochs: keep the call counter in device memory (bumped by a one-thread kernel in front of every GEMM) so calls
        captured into a CUDA graph replay correctly; the partial area is then picked on the device by the counter's parity."""
        self.symm, self.n, self.max_m = symm, n, max_m
        self._epoch_dev = torch.zeros(1, dtype=torch.int32, device=torch.device("cuda", torch.cuda.current_device())) if device_epochs else None
        self.part_bytes = (max_m * n * 2 + 255) // 256 * 256
        self.part_off = [offset, offset + self.part_bytes]
        self.flag_off = offset + 2 * self.part_bytes
        self.n_flags = ((max_m + 127) // 128) * ((n + 255) // 256)
        self.nbytes = 2 * self.part_bytes + self.n_flags * 4
        assert offset + self.nbytes <= symm.nbytes, "symmetric buffer too small for FusedGemmAllReduce"
        symm.view(self.flag_off, (self.n_flags, ), torch.int32).zero_()
        self.epoch = 0

    @staticmethod
    def bytes_needed(max_m, n):
        part = (max_m * n * 2 + 255) // 256 * 256
        return 2 * part + ((max_m + 127) // 128) * ((n + 255) // 256) * 4

    def __call__(self, x, w_shard, residual=None, out=None, max_ctas=0, rows=None):
        """x: [M, K_local] bf16, w_shard: [N, K_local] bf16 -> sum over ranks of x @ w_shard^T (+ residual), [M, N] bf16.
        rows = (lo, hi): reduce-scatter form -- this rank reduces and returns only rows [lo, hi) of the sum ([hi - lo, N]; `residual`
        has that shape too); the ranks' row ranges must tile [0, M)."""
        M, K = x.shape
        N = w_shard.shape[0]
        assert N == self.n and M <= self.max_m and w_shard.shape[1] == K and x.stride(1) == 1 and w_shard.stride(1) == 1
        lo, hi = (0, M) if rows is None else rows
        if out is None:
            out = torch.empty(hi - lo, N, dtype=torch.bfloat16, device=x.device)
        assert out.shape[0] == hi - lo and (residual is None or residual.shape[0] == hi - lo)
        s = self.symm
        arr = (ctypes.c_void_p * s.world)(*s.ptrs)
        if self._epoch_dev is not None:
            rc = native.lib().dsb200_gemm_ar_bf16_dev(native.ptr(x), native.ptr(w_shard), ctypes.cast(arr, ctypes.c_void_p), self.part_off[0],
                                                      self.part_off[1], self.flag_off, native.ptr(self._epoch_dev), native.ptr(out),
                                                      native.ptr(residual), M, N, K, x.stride(0), w_shard.stride(0), out.stride(0),
                                                      residual.stride(0) if residual is not None else 0, s.rank, s.world, int(lo), int(hi),
                                                      int(max_ctas), native.stream_ptr())
            native.check(rc)
            return out
        self.epoch += 1
        common = (native.ptr(x), native.ptr(w_shard), ctypes.cast(arr, ctypes.c_void_p), self.part_off[self.epoch & 1], self.flag_off, self.epoch,
                  native.ptr(out), native.ptr(residual), M, N, K, x.stride(0), w_shard.stride(0), out.stride(0),
                  residual.stride(0) if residual is not None else 0, s.rank, s.world)
        if rows is None:
            rc = native.lib().dsb200_gemm_ar_bf16(*common, int(max_ctas), native.stream_ptr())
        else:
            rc = native.lib().dsb200_gemm_rs_bf16(*common, int(lo), int(hi), int(max_ctas), native.stream_ptr())
        native.check(rc)
        return out


class FusedGemmReduceScatter(FusedGemmAllReduce):
    """Row-parallel linear whose output stays sharded over tokens (sequence-parallel tensor parallelism): rank r gets rows
    [r * M / W, (r + 1) * M / W) of the reduced product.  Same kernel as the all-reduce form; each rank pulls 1 / W of the tiles."""

    def __call__(self, x, w_shard, residual=None, out=None, max_ctas=0):
        M, W, r = x.shape[0], self.symm.world, self.symm.rank
        assert M % W == 0, f"rows ({M}) must divide evenly over the {W} ranks"
        return super().__call__(x, w_shard, residual=residual, out=out, max_ctas=max_ctas, rows=(r * M // W, (r + 1) * M // W))
