"""KV block allocator (free list).  Parity: inference/v2/ragged/blocked_allocator.py:11-105 -- the reference keeps the
linked list in a torch tensor and walks it in Python; here it is the C++ allocator of csrc/host/ragged_host.cpp."""
import ctypes

import torch

from ....ops import native


def _lib():
    l = native.host_lib()
    if not getattr(l, "_ragged_sigs", False):
        P, I = ctypes.c_void_p, ctypes.c_int
        l.dsbh_allocator_create.argtypes = [I]
        l.dsbh_allocator_create.restype = P
        l.dsbh_allocator_destroy.argtypes = [P]
        l.dsbh_allocator_free_blocks.argtypes = [P]
        l.dsbh_allocator_allocate.argtypes = [P, I, P]
        l.dsbh_allocator_free.argtypes = [P, P, I]
        l.dsbh_build_atoms.argtypes = [P, P, P, I, I, I, P, I]
        l.dsbh_pack_batch.argtypes = [P, P, I, P, P, P]
        l._ragged_sigs = True
    return l


class BlockedAllocator:

    def __init__(self, num_blocks: int) -> None:
        if num_blocks < 1:
            raise ValueError(f"Blocked KV-cache must have at least 1 block, provided {num_blocks}")
        self._num_blocks = num_blocks
        self._l = _lib()
        self._h = self._l.dsbh_allocator_create(num_blocks)

    def __del__(self):
        try:
            self._l.dsbh_allocator_destroy(self._h)
        except Exception:
            pass

    def allocate(self, num_blocks: int) -> torch.Tensor:
        out = torch.empty(num_blocks, dtype=torch.int32)
        if self._l.dsbh_allocator_allocate(self._h, num_blocks, out.data_ptr()) != 0:
            raise ValueError(f"Not enough free blocks in the KV-cache to allocate {num_blocks} blocks")
        return out

    def free(self, blocks) -> None:
        if isinstance(blocks, int):
            blocks = [blocks]
        b = torch.as_tensor(blocks, dtype=torch.int32).contiguous()
        if self._l.dsbh_allocator_free(self._h, b.data_ptr(), b.numel()) != 0:
            raise ValueError(f"Invalid or already-free block in {b.tolist()}")

    @property
    def free_blocks(self) -> int:
        return self._l.dsbh_allocator_free_blocks(self._h)
