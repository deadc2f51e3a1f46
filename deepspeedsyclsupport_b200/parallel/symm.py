"""Symmetric (peer-mapped) device memory over CUDA IPC + the P2P collective kernels of csrc/comm/symm.cu.

`SymmetricBuffer(group, nbytes)` gives every rank of `group` a buffer of the same size; each rank maps all peers'
buffers, so kernels can load/store remote HBM directly over NVLink/NVSwitch.  Layout of a buffer:
    [ data region : nbytes ][ flag pad : 4 KiB ]
Cross-rank ordering uses `barrier()` -- a one-CTA kernel on the current stream that publishes an epoch to every peer's
flag pad and waits for all peers (bounded spin) -- so no host synchronisation is involved.

`EmulatedPeers` builds the same structure inside ONE process on ONE GPU (W local buffers); it is how the collective
kernels are tested on a single GPU (each "rank's" kernel is launched in turn, no cross-kernel waiting).
"""
import ctypes

import torch

from .. import comm as dist
from ..ops import native

native.register_sigs({
    "dsb200_symm_alloc": "l pp",
    "dsb200_symm_open": "pp",
    "dsb200_symm_close": "p",
    "dsb200_symm_free": "p",
    "dsb200_symm_barrier": "p l ii i p",
    "dsb200_symm_barrier_dev": "p l ii p p",
    "dsb200_p2p_allgather": "p l p l iii p",
    "dsb200_p2p_alltoall": "p l p l iii p",
    "dsb200_p2p_reduce_scatter": "p l i p i l ii f i p i p",
    "dsb200_p2p_rs_adam": "p l pppp l ii f fffff iii i p",
    "dsb200_p2p_allreduce_bf16": "p l p l iii p",
    "dsb200_p2p_allgather_ce": "p l p l ii p",
    "dsb200_p2p_fp8_reduce_scatter": "p ll p i l ii f i p i p",
    "dsb200_p2p_rs_adam_range": "pp lll l pppppp ii f fffff iii i p",
    "dsb200_nvls_reduce_scatter": "p l i p i l ii f i p i p",
    "dsb200_nvls_allreduce_bf16": "p l l ii i p",
    "dsb200_nvls_allgather": "p l p l i i p",
})

FLAG_PAD = 4096
_DT = {torch.bfloat16: 0, torch.float16: 1, torch.float32: 2}


class _RawCudaArray:
    """Minimal __cuda_array_interface__ carrier so torch can alias raw device memory without copying."""

    def __init__(self, ptr, nbytes):
        self.__cuda_array_interface__ = {"shape": (nbytes, ), "typestr": "|u1", "data": (ptr, False), "version": 3}


def os_env_flag(name):
    import os
    return os.environ.get(name, "0") == "1"


def _alias(ptr, nbytes, device):
    t = torch.as_tensor(_RawCudaArray(ptr, nbytes), device=device)
    return t


class _PeerOps:
    """Kernels shared by the real and the emulated flavour.  `self.ptrs` = list of W base pointers, `self.rank`."""

    def _table(self):
        arr = (ctypes.c_void_p * self.world)(*self.ptrs)
        return arr

    def view(self, offset, shape, dtype, rank=None):
        """Tensor aliasing [offset, offset + numel*itemsize) of (this rank's, or `rank`'s mapped) buffer."""
        n = 1
        for s in shape:
            n *= s
        nbytes = n * torch.empty(0, dtype=dtype).element_size()
        assert offset + nbytes <= self.nbytes, "symmetric buffer overflow"
        base = self.ptrs[self.rank if rank is None else rank]
        key = (base, offset, nbytes, dtype)
        flat = self._views.get(key)
        if flat is None:
            flat = self._views[key] = _alias(base + offset, nbytes, self.device).view(dtype)
        return flat.view(*shape)

    mc_ptr = 0  # multicast (NVLS) mapping of the buffer; 0 = not available

    def all_gather_ce(self, out, src_offset, bytes_per_rank, rank=None):
        """All-gather on the copy engines (peer -> local DMA per rank, no SMs): the form used under compute."""
        r = self.rank if rank is None else rank
        rc = native.lib().dsb200_p2p_allgather_ce(ctypes.cast(self._table(), ctypes.c_void_p), src_offset, native.ptr(out),
                                                  bytes_per_rank, r, self.world, native.stream_ptr())
        native.check(rc)
        return out

    def nvls_reduce_scatter(self, dst, src_offset, src_dtype, n_per_rank, scale=1.0, accumulate=False, sumsq=None, max_ctas=0):
        """Reduce-scatter through the NVSwitch: `multimem.ld_reduce` over the multicast mapping sums slice `rank` of every
        rank's buffer inside the switch, so this GPU receives n elements instead of reading (W-1)*n from peers."""
        assert self.mc_ptr, "no multicast mapping for this buffer"
        rc = native.lib().dsb200_nvls_reduce_scatter(self.mc_ptr, src_offset, _DT[src_dtype], native.ptr(dst), _DT[dst.dtype],
                                                     n_per_rank, self.rank, self.world, float(scale), int(accumulate),
                                                     native.ptr(sumsq), max_ctas, native.stream_ptr())
        native.check(rc)
        return dst

    def nvls_all_reduce_bf16(self, offset, n, max_ctas=0):
        """In-place bf16 all-reduce of the n elements at `offset` of every rank's buffer (ld_reduce + multicast store)."""
        assert self.mc_ptr, "no multicast mapping for this buffer"
        self.barrier()  # every rank's addend is written
        rc = native.lib().dsb200_nvls_allreduce_bf16(self.mc_ptr, offset, n, self.rank, self.world, max_ctas, native.stream_ptr())
        native.check(rc)
        self.barrier()  # every slice's sums have landed everywhere

    def nvls_all_gather(self, dst_offset, src, max_ctas=0):
        """Push all-gather: `src` (local, contiguous) is multicast-stored into every rank's buffer at dst_offset + rank*nbytes.
        The caller runs `barrier()` before readers consume it."""
        assert self.mc_ptr, "no multicast mapping for this buffer"
        nbytes = src.numel() * src.element_size()
        rc = native.lib().dsb200_nvls_allgather(self.mc_ptr, dst_offset, native.ptr(src), nbytes, self.rank, max_ctas,
                                                native.stream_ptr())
        native.check(rc)

    def all_gather(self, out, src_offset, bytes_per_rank, rank=None, max_ctas=0):
        r = self.rank if rank is None else rank
        rc = native.lib().dsb200_p2p_allgather(ctypes.cast(self._table(), ctypes.c_void_p), src_offset, native.ptr(out),
                                               bytes_per_rank, r, self.world, max_ctas, native.stream_ptr())
        native.check(rc)
        return out

    def all_to_all(self, out, src_offset, bytes_per_chunk, rank=None, max_ctas=0):
        """Equal-split all_to_all_single: out[p] = chunk `rank` of peer p's buffer (the caller has written its `world` send
        chunks at `src_offset` and every rank has passed a barrier since)."""
        r = self.rank if rank is None else rank
        rc = native.lib().dsb200_p2p_alltoall(ctypes.cast(self._table(), ctypes.c_void_p), src_offset, native.ptr(out),
                                              bytes_per_chunk, r, self.world, max_ctas, native.stream_ptr())
        native.check(rc)
        return out

    def reduce_scatter(self, dst, src_offset, src_dtype, n_per_rank, scale=1.0, accumulate=False, sumsq=None, rank=None,
                       max_ctas=0):
        r = self.rank if rank is None else rank
        rc = native.lib().dsb200_p2p_reduce_scatter(ctypes.cast(self._table(), ctypes.c_void_p), src_offset, _DT[src_dtype],
                                                    native.ptr(dst), _DT[dst.dtype], n_per_rank, r, self.world, float(scale),
                                                    int(accumulate), native.ptr(sumsq), max_ctas, native.stream_ptr())
        native.check(rc)
        return dst

    def rs_adam(self, src_offset, p, m, v, lp, n_per_rank, *, scale, lr, beta1, beta2, eps, weight_decay, adamw_mode=True,
                bias_correction=True, step=1, rank=None, max_ctas=0):
        r = self.rank if rank is None else rank
        rc = native.lib().dsb200_p2p_rs_adam(ctypes.cast(self._table(), ctypes.c_void_p), src_offset, native.ptr(p), native.ptr(m),
                                             native.ptr(v), native.ptr(lp), n_per_rank, r, self.world, float(scale), lr, beta1, beta2,
                                             eps, weight_decay, int(adamw_mode), int(bias_correction), int(step), max_ctas,
                                             native.stream_ptr())
        native.check(rc)

    def fp8_reduce_scatter(self, dst, q_offset, s_offset, n_per_rank, scale=1.0, accumulate=False, sumsq=None, max_ctas=0):
        """qgZ consumer: every rank holds its W*n gradient buffer as block-scaled fp8 (payload at q_offset, fp32 scale per
        128 elements at s_offset); this rank pulls slice `rank` from all peers, dequantises, sums, scales into `dst`."""
        rc = native.lib().dsb200_p2p_fp8_reduce_scatter(ctypes.cast(self._table(), ctypes.c_void_p), q_offset, s_offset,
                                                        native.ptr(dst), _DT[dst.dtype], n_per_rank, self.rank, self.world,
                                                        float(scale), int(accumulate), native.ptr(sumsq), max_ctas,
                                                        native.stream_ptr())
        native.check(rc)
        return dst

    def rs_adam_range(self, src_offset, n_per_rank, elem_lo, p, m, v, lp, *, scale, lr, beta1, beta2, eps, weight_decay,
                      adamw_mode=True, bias_correction=True, step=1, g_out=None, sumsq=None, nvls=False, max_ctas=0):
        """Reduce-scatter fused with AdamW on elements [elem_lo, elem_lo + p.numel()) of this rank's slice of the bf16 gradient
        buffer at `src_offset` (W * n_per_rank elements).  nvls=True takes the sum from the switch.  `g_out` (bf16) keeps the
        reduced gradient, `sumsq` accumulates its squared norm."""
        rc = native.lib().dsb200_p2p_rs_adam_range(ctypes.cast(self._table(), ctypes.c_void_p), self.mc_ptr if nvls else None,
                                                   src_offset, n_per_rank, elem_lo, p.numel(), native.ptr(p), native.ptr(m),
                                                   native.ptr(v), native.ptr(lp), native.ptr(g_out), native.ptr(sumsq), self.rank,
                                                   self.world, float(scale), lr, beta1, beta2, eps, weight_decay, int(adamw_mode),
                                                   int(bias_correction), int(step), max_ctas, native.stream_ptr())
        native.check(rc)

    def all_reduce_bf16(self, dst, src_offset, n, rank=None, max_ctas=0, scratch_offset=None, algo="auto"):
        """dst[n] = sum over ranks of the bf16 vector at `src_offset` of every rank's buffer.
        one-shot : every rank reads all W vectors and reduces locally (W-1)*n remote reads, one launch: lowest latency,
                   right for decode-sized messages.
        two-shot : reduce-scatter into a symmetric scratch slice, device barrier, all-gather of the slices:
                   2*(W-1)/W*n remote reads: right for large messages on many GPUs.  Needs `scratch_offset` (n/W
                   bf16 elements of this symmetric buffer) and a real multi-process buffer (it uses the device barrier).
        auto     : two-shot when a scratch region is given, W > 2 and the message is >= 4 MiB."""
        r = self.rank if rank is None else rank
        can_two = scratch_offset is not None and rank is None and hasattr(self, "barrier") and n % (8 * self.world) == 0
        if algo == "two_shot":
            assert can_two, "two-shot all-reduce needs scratch_offset, a real SymmetricBuffer and n % (8*world) == 0"
        if (algo == "two_shot") or (algo == "auto" and can_two and self.world > 2 and n * 2 >= (4 << 20)):
            per = n // self.world
            mine = self.view(scratch_offset, (per, ), torch.bfloat16)
            self.reduce_scatter(mine, src_offset, torch.bfloat16, per, scale=1.0, max_ctas=max_ctas)
            self.barrier()  # every rank's reduced slice is complete before anyone gathers it
            return self.all_gather(dst, scratch_offset, per * 2, max_ctas=max_ctas)
        rc = native.lib().dsb200_p2p_allreduce_bf16(ctypes.cast(self._table(), ctypes.c_void_p), src_offset, native.ptr(dst), n, r,
                                                    self.world, max_ctas, native.stream_ptr())
        native.check(rc)
        return dst


class SymmetricBuffer(_PeerOps):
    """backend: "torch" = torch.distributed._symmetric_memory (CUDA VMM allocation, also bound to an NVSwitch multicast
    object when the fabric supports it -> `mc_ptr` for the NVLS kernels); "ipc" = cudaMalloc + CUDA-IPC handles (P2P only).
    Default: DSB200_SYMM_BACKEND, else "torch" with a fall-back to "ipc"."""

    def __init__(self, group, nbytes, backend=None):
        import os
        assert torch.cuda.is_available()
        self.group = group
        self.world = dist.get_world_size(group)
        self.rank = dist.get_rank(group)
        self.nbytes = (int(nbytes) + 255) // 256 * 256
        self.device = torch.device("cuda", torch.cuda.current_device())
        self._views = {}
        self._epoch = 0
        self._opened = []
        self.local_ptr = None
        self.mc_ptr = 0
        backend = backend or os.environ.get("DSB200_SYMM_BACKEND", "torch")
        self.backend = None
        if backend == "torch":
            ok = 0
            try:
                self._init_torch(group)
                ok = 1
            except Exception as e:  # allocator / rendezvous not available on this stack
                self._torch_error = f"{type(e).__name__}: {e}"
            # the choice is collective: one rank without it moves every rank to the IPC flavour
            flag = torch.tensor([ok], dtype=torch.int32, device=self.device)
            dist.all_reduce(flag, op=dist.ReduceOp.MIN, group=group)
            if int(flag.item()) == 1:
                self.backend = "torch"
            else:
                self._symm_tensor = self._symm_handle = None
                self.mc_ptr = 0
        if self.backend is None:
            self._init_ipc(group)
            self.backend = "ipc"
        dist.barrier(group=group)

    def _init_torch(self, group):
        import torch.distributed as tdist
        import torch.distributed._symmetric_memory as symm_mem
        pg = group if group is not None else tdist.group.WORLD
        t = symm_mem.empty(self.nbytes + FLAG_PAD, dtype=torch.uint8, device=self.device)
        t.zero_()
        torch.cuda.synchronize()
        hdl = symm_mem.rendezvous(t, pg)
        self._symm_tensor, self._symm_handle = t, hdl
        off = int(getattr(hdl, "offset", 0) or 0)  # the tensor may sit inside a larger symmetric allocation
        self.ptrs = [int(p) + off for p in hdl.buffer_ptrs]
        if self.ptrs[self.rank] != t.data_ptr():
            off = t.data_ptr() - int(hdl.buffer_ptrs[self.rank])
            self.ptrs = [int(p) + off for p in hdl.buffer_ptrs]
        assert len(self.ptrs) == self.world and 0 <= off and off % 256 == 0
        self.local_ptr = None
        mc = 0
        if not os_env_flag("DSB200_DISABLE_NVLS"):
            try:
                mc = int(hdl.multicast_ptr or 0)
                if mc:
                    mc += off
            except Exception:
                mc = 0
        self.mc_ptr = mc

    def _init_ipc(self, group):
        lib = native.lib()
        ptr = ctypes.c_void_p()
        handle = (ctypes.c_ubyte * 64)()
        rc = lib.dsb200_symm_alloc(self.nbytes + FLAG_PAD, ctypes.addressof(ptr), ctypes.addressof(handle))
        native.check(rc, 0)
        self.local_ptr = ptr.value
        handles = [None] * self.world
        dist.all_gather_object(handles, bytes(handle), group=group)
        self.ptrs = []
        for r, h in enumerate(handles):
            if r == self.rank:
                self.ptrs.append(self.local_ptr)
                continue
            hp = (ctypes.c_ubyte * 64).from_buffer_copy(h)
            p = ctypes.c_void_p()
            rc = lib.dsb200_symm_open(ctypes.addressof(hp), ctypes.addressof(p))
            native.check(rc, 0)
            self.ptrs.append(p.value)
            self._opened.append(p.value)

    def use_device_epochs(self):
        """Switch `barrier()` to the graph-capturable form: the epoch lives in a device counter that the barrier kernel advances
        itself, so a barrier captured into a CUDA graph replays correctly.  For buffers whose barriers are all issued on ONE
        stream in the same order on every rank (the FastGen tensor-parallel collectives); call it collectively before the
        first use that must be capturable."""
        if getattr(self, "_epoch_dev", None) is None:
            self._epoch_dev = torch.zeros(1, dtype=torch.int32, device=self.device)
        return self

    def barrier(self):
        """Device-side barrier across the group on the current stream (no host sync)."""
        if getattr(self, "_epoch_dev", None) is not None:
            # own flag pad (second half of the 4 KiB pad): device epochs and host epochs never mix on one set of flags
            rc = native.lib().dsb200_symm_barrier_dev(ctypes.cast(self._table(), ctypes.c_void_p), self.nbytes + FLAG_PAD // 2, self.rank,
                                                      self.world, native.ptr(self._epoch_dev), native.stream_ptr())
            native.check(rc)
            return
        self._epoch += 1
        rc = native.lib().dsb200_symm_barrier(ctypes.cast(self._table(), ctypes.c_void_p), self.nbytes, self.rank, self.world,
                                              self._epoch, native.stream_ptr())
        native.check(rc)

    def close(self):
        lib = native.lib()
        torch.cuda.synchronize()
        for p in self._opened:
            lib.dsb200_symm_close(p)
        self._opened = []
        if self.local_ptr:
            lib.dsb200_symm_free(self.local_ptr)
            self.local_ptr = None
        self._views = {}
        self._symm_tensor = self._symm_handle = None


class EmulatedPeers(_PeerOps):
    """W peer buffers inside one process (single-GPU tests of the P2P kernels)."""

    def __init__(self, world, nbytes):
        self.world = world
        self.rank = 0
        self.nbytes = (int(nbytes) + 255) // 256 * 256
        self.device = torch.device("cuda", torch.cuda.current_device())
        self._views = {}
        self._bufs = [torch.zeros(self.nbytes + FLAG_PAD, dtype=torch.uint8, device=self.device) for _ in range(world)]
        self.ptrs = [b.data_ptr() for b in self._bufs]


class PeerAllToAll(object):
    """Equal-split all_to_all_single over NVLink peer memory for one process group: the send tensor is copied into a symmetric
    staging buffer, a device barrier publishes it, every rank pulls its chunks from the peers (csrc/comm/symm.cu
    alltoall_pull_kernel), a second barrier frees the staging buffer.  No NCCL call, no host synchronisation.  The staging
    buffer grows on demand (collective: every rank of the group sees the same sizes)."""
    _instances = {}

    def __init__(self, group, nbytes):
        self.group = group
        self.buf = SymmetricBuffer(group, nbytes)

    @classmethod
    def for_group(cls, group, nbytes):
        key = tuple(dist.get_all_ranks_from_group(group)) if group is not None and hasattr(dist, "get_all_ranks_from_group") else group
        inst = cls._instances.get(key)
        if inst is None or inst.buf.nbytes < nbytes:
            if inst is not None:
                inst.buf.close()
            inst = cls(group, max(int(nbytes), 1 << 22))
            cls._instances[key] = inst
        return inst

    @staticmethod
    def usable(tensor, group):
        import os
        if not (tensor.is_cuda and tensor.is_contiguous() and os.environ.get("DSB200_DISABLE_P2P", "0") != "1"):
            return False
        world = dist.get_world_size(group)
        nbytes = tensor.numel() * tensor.element_size()
        return 1 < world <= 16 and nbytes % (16 * world) == 0 and world <= torch.cuda.device_count()

    def __call__(self, output, input):
        world = self.buf.world
        nbytes = input.numel() * input.element_size()
        stage = self.buf.view(0, (nbytes, ), torch.uint8)
        stage.copy_(input.reshape(-1).view(torch.uint8))
        self.buf.barrier()   # every rank's send chunks are in place
        self.buf.all_to_all(output, 0, nbytes // world)
        self.buf.barrier()   # everyone has pulled: the staging buffer may be overwritten by the next call
        return output
