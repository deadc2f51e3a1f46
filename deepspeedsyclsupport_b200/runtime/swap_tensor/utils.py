"""Swap buffer pool + helpers (parity: deepspeed/runtime/swap_tensor/utils.py:96 SwapBufferPool)."""
import torch

MIN_AIO_BYTES = 1024 ** 2
AIO_ALIGNED_BYTES = 1024


def swap_in_tensors(swap_handle, tensor_buffers, swap_paths):
    for buffer, path in zip(tensor_buffers, swap_paths):
        assert swap_handle.async_pread(buffer, path) == 0


def swap_out_tensors(swap_handle, tensor_buffers, swap_paths):
    for buffer, path in zip(tensor_buffers, swap_paths):
        assert swap_handle.async_pwrite(buffer, path) == 0


def get_sized_buffer(buffer, num_elems):
    assert num_elems <= buffer.numel(), f"num_elems {num_elems} > buffer {buffer.numel()}"
    return buffer.narrow(0, 0, num_elems) if num_elems < buffer.numel() else buffer


class SwapBufferPool:
    """Fixed set of (pinned) host buffers handed out round-robin for in-flight swaps."""

    def __init__(self, buffers):
        assert all(b.device.type == "cpu" for b in buffers)
        self.buffers = list(buffers)
        self.free = list(range(len(buffers)))
        self.used = {}

    def has_space(self, numel=0):
        return bool(self.free) and (numel == 0 or self.buffers[self.free[0]].numel() >= numel)

    def allocate(self, key, numel):
        assert self.has_space(numel), "swap buffer pool exhausted"
        idx = self.free.pop(0)
        self.used[key] = idx
        return get_sized_buffer(self.buffers[idx], numel)

    def release(self, key):
        self.free.append(self.used.pop(key))

    def release_all(self):
        for k in list(self.used):
            self.release(k)
