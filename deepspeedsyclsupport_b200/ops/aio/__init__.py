"""`aio_handle`: asynchronous file reads/writes of tensors for the NVMe tier (csrc/host/aio.cpp).
API parity: deepspeed.ops.aio (AsyncIOBuilder().load().aio_handle) -- csrc/aio/py_lib/py_ds_aio.cpp:13-45:
aio_handle(block_size, queue_depth, single_submit, overlap_events, num_threads) with sync_pread / sync_pwrite /
async_pread / async_pwrite / wait / new_cpu_locked_tensor / free_cpu_locked_tensor and the getters."""
import ctypes

import torch

from .. import native


def _lib():
    l = native.host_lib()
    if not getattr(l, "_aio_sigs", False):
        P, I, L = ctypes.c_void_p, ctypes.c_int, ctypes.c_int64
        l.dsbh_aio_create.argtypes = [L, I, I, I, I]
        l.dsbh_aio_create.restype = P
        l.dsbh_aio_destroy.argtypes = [P]
        l.dsbh_aio_wait.argtypes = [P]
        l.dsbh_aio_wait.restype = L
        l.dsbh_aio_pread.argtypes = [P, P, L, ctypes.c_char_p, I, L]
        l.dsbh_aio_pwrite.argtypes = [P, P, L, ctypes.c_char_p, I, L]
        l.dsbh_aio_block_size.argtypes = [P]
        l.dsbh_aio_block_size.restype = L
        l.dsbh_aio_queue_depth.argtypes = [P]
        l.dsbh_aio_thread_count.argtypes = [P]
        l._aio_sigs = True
    return l


class aio_handle:

    def __init__(self, block_size=1048576, queue_depth=8, single_submit=False, overlap_events=True, num_threads=1):
        self._l = _lib()
        self._h = self._l.dsbh_aio_create(block_size, queue_depth, int(single_submit), int(overlap_events), num_threads)
        self._single_submit, self._overlap_events = single_submit, overlap_events
        self._pending = []  # keep tensors alive while their I/O is in flight

    def __del__(self):
        try:
            if self._h:
                self._l.dsbh_aio_destroy(self._h)
                self._h = None
        except Exception:
            pass

    def get_block_size(self):
        return self._l.dsbh_aio_block_size(self._h)

    def get_queue_depth(self):
        return self._l.dsbh_aio_queue_depth(self._h)

    def get_single_submit(self):
        return self._single_submit

    def get_overlap_events(self):
        return self._overlap_events

    def get_thread_count(self):
        return self._l.dsbh_aio_thread_count(self._h)

    def _io(self, tensor, filename, write, async_op, offset=0):
        assert tensor.device.type == "cpu" and tensor.is_contiguous(), "aio works on contiguous host tensors"
        fn = self._l.dsbh_aio_pwrite if write else self._l.dsbh_aio_pread
        rc = fn(self._h, tensor.data_ptr(), tensor.numel() * tensor.element_size(), str(filename).encode(), int(async_op), offset)
        if rc != 0:
            raise IOError(f"aio {'write' if write else 'read'} of {filename} failed (code {rc})")
        if async_op:
            self._pending.append(tensor)
        return 0 if async_op else 1

    def sync_pread(self, buffer, filename):
        return self._io(buffer, filename, False, False)

    def sync_pwrite(self, buffer, filename):
        return self._io(buffer, filename, True, False)

    def async_pread(self, buffer, filename):
        return self._io(buffer, filename, False, True)

    def async_pwrite(self, buffer, filename):
        return self._io(buffer, filename, True, True)

    def read(self, buffer, filename, validate=False):
        return self.sync_pread(buffer, filename)

    def write(self, buffer, filename, validate=False):
        return self.sync_pwrite(buffer, filename)

    def pread(self, buffer, filename, validate=False, async_op=False):
        return self._io(buffer, filename, False, async_op)

    def pwrite(self, buffer, filename, validate=False, async_op=False):
        return self._io(buffer, filename, True, async_op)

    def wait(self):
        n = self._l.dsbh_aio_wait(self._h)
        self._pending.clear()
        if n < 0:
            raise IOError("asynchronous I/O failed")
        return int(n)

    def new_cpu_locked_tensor(self, num_elem, example_tensor):
        t = torch.empty(num_elem, dtype=example_tensor.dtype)
        return t.pin_memory() if torch.cuda.is_available() else t

    def free_cpu_locked_tensor(self, tensor):
        return True


class _AioModule:
    aio_handle = aio_handle

    @staticmethod
    def aio_read(buffer, filename, block_size, queue_depth, single_submit, overlap_events, validate):
        return aio_handle(block_size, queue_depth, single_submit, overlap_events, 1).sync_pread(buffer, filename)

    @staticmethod
    def aio_write(buffer, filename, block_size, queue_depth, single_submit, overlap_events, validate):
        return aio_handle(block_size, queue_depth, single_submit, overlap_events, 1).sync_pwrite(buffer, filename)


class AsyncIOBuilder:
    NAME = "async_io"

    def is_compatible(self, verbose=True):
        return True

    def load(self, verbose=True):
        return _AioModule
