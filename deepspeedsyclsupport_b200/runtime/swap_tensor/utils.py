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


def get_sized_buffers(buffer_list, num_elems_list):
    """Prefix views of several buffers at once (parity: swap_tensor/utils.py:236-241)."""
    return [get_sized_buffer(b, n) for b, n in zip(buffer_list, num_elems_list)]


class SwapBuffer:
    """One host staging buffer packed with several tensors.  Each tensor gets an I/O-aligned slot (``swap`` view,
    the bytes that travel to the file) and a ``compute`` view of its true length (parity: swap_tensor/utils.py:36-93)."""

    def __init__(self, buffer):
        self.buffer = buffer
        self.reset()

    def reset(self):
        self.offset = 0
        self.swap_tensors = {}
        self.compute_tensors = {}
        self.swap_paths = {}
        self.num_elem = 0

    def has_space(self, numel):
        return self.offset + numel <= self.buffer.numel()

    def allocate_tensor(self, swap_path, numel, aligned_numel):
        assert self.has_space(aligned_numel), f"swap buffer full: need {aligned_numel} at offset {self.offset} of {self.buffer.numel()}"
        assert self.offset not in self.swap_tensors
        slot = self.offset
        self.swap_tensors[slot] = self.buffer.narrow(0, slot, aligned_numel)
        self.compute_tensors[slot] = self.buffer.narrow(0, slot, numel)
        self.swap_paths[slot] = swap_path
        self.offset += aligned_numel
        self.num_elem += numel
        return self.swap_tensors[slot], self.compute_tensors[slot]

    def insert_tensor(self, tensor, swap_path, aligned_numel):
        swap_tensor, compute_tensor = self.allocate_tensor(swap_path, tensor.numel(), aligned_numel)
        compute_tensor.data.copy_(tensor.data.reshape(-1))
        return swap_tensor, compute_tensor

    def get_swap_tensors(self):
        return list(self.swap_tensors.values())

    def get_swap_paths(self):
        return list(self.swap_paths.values())

    def get_compute_tensors(self):
        return list(self.compute_tensors.values())

    def get_num_elem(self):
        return self.num_elem

    def get_swap_tensor(self, offset):
        return self.swap_tensors.get(offset)

    def get_compute_tensor(self, offset):
        return self.compute_tensors.get(offset)

    def get_swap_path(self, offset):
        return self.swap_paths.get(offset)


class SwapBufferManager:
    """Owns ``count`` pinned host buffers of ``num_elems`` elements and lends them out
    (parity: swap_tensor/utils.py:180-227).  Pinning is skipped on a host-only build."""

    def __init__(self, num_elems, count, dtype):
        self.num_elems = num_elems
        self.count = count
        self.dtype = dtype
        pin = torch.cuda.is_available()
        self.all_buffers = [torch.zeros(num_elems, dtype=dtype, device="cpu", pin_memory=pin) for _ in range(count)]
        self.free_buffer_index = list(range(count))
        self.used_buffer_index = {}
        self.gigabytes = self.all_buffers[0].element_size() * num_elems * count / (1024 ** 3)

    def allocate(self, num_elems, count, dtype):
        assert dtype == self.dtype, f"manager holds {self.dtype}, asked for {dtype}"
        assert num_elems <= self.num_elems, f"asked for {num_elems} elements, buffers hold {self.num_elems}"
        if count > len(self.free_buffer_index):
            return None
        taken, self.free_buffer_index = self.free_buffer_index[-count:], self.free_buffer_index[:-count]
        out = []
        for i in taken:
            view = self.all_buffers[i].narrow(0, 0, num_elems)
            self.used_buffer_index[id(view)] = i
            out.append(view)
        return out

    def allocate_all(self, num_elems, dtype):
        return self.allocate(num_elems=num_elems, count=len(self.free_buffer_index), dtype=dtype)

    def free(self, buffers):
        for b in buffers:
            assert id(b) in self.used_buffer_index, "buffer was not handed out by this manager"
        for b in buffers:
            self.free_buffer_index.append(self.used_buffer_index.pop(id(b)))
