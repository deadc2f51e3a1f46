"""Error-compensated 1-bit all-reduce (the collective under OnebitAdam / OnebitLamb / ZeroOneAdam).

Algorithm (parity: deepspeed/runtime/comm/nccl.py:52-170, mpi.py, compressed.py):
  worker: v = x + e_w ; s_w = ||v||/sqrt(n) ; e_w = v - s_w*sign(v) ; send sign bits of chunk j to rank j (all-to-all) and
          s_w to everybody
  server: m = mean_w(s_w * sign_w(chunk)) + e_s ; s_s = ||m||/sqrt(c) ; e_s = m - s_s*sign(m) ; all-gather bits + s_s
  result: x = concat_r(s_s[r] * sign_r)
B200 design: the pack/unpack + error-feedback arithmetic are two fused CUDA kernels per direction (csrc/quant/onebit.cu)
instead of cupy packbits and ~10 elementwise torch launches; the wire format is 1 bit/element over NCCL all-to-all /
all-gather.  On CPU tensors (gloo tests) the same math runs in torch.
"""
import numpy as np
import torch

from ... import comm as dist
from ...ops import native

native.register_sigs({
    "dsb200_onebit_compress": "pppp l p",
    "dsb200_onebit_decompress": "ppp i l i p",
})

_BITS = torch.tensor([128, 64, 32, 16, 8, 4, 2, 1], dtype=torch.uint8)


def onebit_compress(x, err):
    """Returns (packed uint8 [n/8], scale fp32 scalar tensor); updates `err` in place (x is left untouched)."""
    n = x.numel()
    assert n % 8 == 0 and err.numel() == n
    if x.is_cuda:
        packed = torch.empty(n // 8, dtype=torch.uint8, device=x.device)
        scale_io = torch.empty(2, dtype=torch.float32, device=x.device)
        native.check(native.lib().dsb200_onebit_compress(native.ptr(x), native.ptr(err), native.ptr(packed), native.ptr(scale_io), n,
                                                         native.stream_ptr()), 2)
        return packed, scale_io[:1]
    v = x + err
    scale = (torch.linalg.norm(v) / np.sqrt(n)).reshape(1)
    pos = v >= 0
    err.copy_(v - scale * (pos.float() * 2 - 1))
    packed = (pos.view(-1, 8).to(torch.uint8) * _BITS).sum(1).to(torch.uint8)
    return packed, scale


def onebit_decompress(packed, scales, mode):
    """packed [rows, chunk_bytes]; mode 0 -> mean over rows [chunk_bytes*8]; mode 1 -> expanded [rows*chunk_bytes*8]."""
    rows, cb = packed.shape
    if packed.is_cuda:
        out = torch.empty(cb * 8 if mode == 0 else rows * cb * 8, dtype=torch.float32, device=packed.device)
        native.check(native.lib().dsb200_onebit_decompress(native.ptr(packed), native.ptr(scales), native.ptr(out), rows, cb, mode,
                                                           native.stream_ptr()))
        return out
    bits = ((packed.unsqueeze(-1) & _BITS) != 0).float() * 2 - 1  # [rows, cb, 8]
    signed = bits.view(rows, cb * 8) * scales.view(rows, 1)
    return signed.mean(0) if mode == 0 else signed.reshape(-1)


class CompressedBackend(object):

    def __init__(self, mpu=None):
        if mpu is None:
            self.world_group = dist.new_group(ranks=list(range(dist.get_world_size())))
        else:
            self.mpu = mpu
            self.world_group = self.mpu.get_data_parallel_group()
        self.rank = dist.get_rank(group=self.world_group)
        self.size = dist.get_world_size(group=self.world_group)

    def compressed_allreduce(self, buffer_m: torch.Tensor, worker_error, server_error, local_rank=0):
        original_shape = buffer_m.size()
        if len(original_shape) > 1:
            buffer_m = torch.flatten(buffer_m)
        original_size = buffer_m.numel()
        worker_error_size = worker_error.numel()
        if original_size != worker_error_size:
            work = torch.zeros(worker_error_size, dtype=torch.float32, device=buffer_m.device)
            work[:original_size].copy_(buffer_m)
        else:
            work = buffer_m.float() if buffer_m.dtype != torch.float32 else buffer_m
        packed, worker_scale = onebit_compress(work, worker_error)
        chunk_bytes = worker_error_size // 8 // self.size
        recv_sign = torch.empty(self.size, chunk_bytes, dtype=torch.uint8, device=packed.device)
        dist.all_to_all_single(recv_sign.view(-1), packed, group=self.world_group)
        recv_scale = torch.empty(self.size, dtype=torch.float32, device=packed.device)
        dist.all_gather_into_tensor(recv_scale, worker_scale, group=self.world_group)
        server_m = onebit_decompress(recv_sign, recv_scale, mode=0)
        server_packed, server_scale = onebit_compress(server_m, server_error)
        all_sign = torch.empty(self.size, chunk_bytes, dtype=torch.uint8, device=packed.device)
        dist.all_gather_into_tensor(all_sign.view(-1), server_packed, group=self.world_group)
        all_scale = torch.empty(self.size, dtype=torch.float32, device=packed.device)
        dist.all_gather_into_tensor(all_scale, server_scale, group=self.world_group)
        full = onebit_decompress(all_sign, all_scale, mode=1)
        buffer_m.copy_(full[:original_size])
        if len(original_shape) > 1:
            buffer_m = buffer_m.reshape(original_shape)
        return buffer_m
