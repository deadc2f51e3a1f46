"""A sub-allocator over one pre-allocated flat buffer, with compaction (parity: deepspeed/runtime/zero/
contiguous_memory_allocator.py:16-287).

Used where many short-lived, differently sized tensors would otherwise fragment the caching allocator (gathered parameters of
offloaded models, activation-checkpoint partitions).  Tensors handed out are views of the buffer; when no gap is large enough
but the total free space is, live tensors are slid towards the front (their `.data`, and the `.data` of parameters assigned to
them, are re-pointed), so callers keep using the same tensor objects."""
import torch

from ...utils.logging import log_dist


def print_rank_0(message):
    log_dist(message, ranks=[0])


class ContiguousMemoryAllocator(object):

    def __init__(self, size, dtype, device):
        self.buffer = torch.zeros(size, dtype=dtype, device=device)
        self.size = size
        self.dtype = dtype
        self.total_free = size
        self.max_allocated = 0
        self.count = 0
        # live blocks: tensor id -> [address, numel];  handed-out tensors and the parameters bound to them, by tensor id
        self.tensor_addresses = {}
        self.tensor_sizes = {}
        self.tensor_map = {}
        self.id_to_params = {}

    # ---- bookkeeping
    def _live_sorted(self):
        return sorted(((addr, self.tensor_sizes[tid], tid) for tid, addr in self.tensor_addresses.items()))

    def _gaps(self):
        """[(address, size)] of free ranges, in address order."""
        gaps, cursor = [], 0
        for addr, size, _ in self._live_sorted():
            if addr > cursor:
                gaps.append((cursor, addr - cursor))
            cursor = addr + size
        if cursor < self.size:
            gaps.append((cursor, self.size - cursor))
        return gaps

    @property
    def contiguous_sizes(self):
        """address -> size of every free range (the reference's public bookkeeping)."""
        return dict(self._gaps())

    def _largest_contiguous(self):
        return max((s for _, s in self._gaps()), default=0)

    # ---- API
    def allocate_tensor(self, size):
        """A zero-copy view of `size` elements.  Compacts first when fragmentation is the only obstacle."""
        assert size <= self.total_free, f"Not enough memory in buffer. Allocated {self.size - self.total_free} of {self.size}, requested {size}"
        if self._largest_contiguous() < size:
            print_rank_0("Needs defragmentation to allocate. Before Defragmentation:")
            self.print_allocation(resolution=100)
            self._defragment_memory()
            print_rank_0("After defragmentation:")
            self.print_allocation(resolution=100)
        # best fit: the smallest gap that holds the request
        addr = min(((s, a) for a, s in self._gaps() if s >= size))[1]
        self.total_free -= size
        self.max_allocated = max(self.max_allocated, self.size - self.total_free)
        t = self.buffer.narrow(0, addr, size)
        tid = id(t)
        self.tensor_addresses[tid], self.tensor_sizes[tid], self.tensor_map[tid] = addr, size, t
        self.id_to_params[tid] = []
        self.count += 1
        return t

    def assign_to_param(self, tensor, param, numel, shape):
        """param.data becomes the first `numel` elements of `tensor`, viewed as `shape`; follows the tensor through compaction."""
        tid = id(tensor)
        assert tid in self.tensor_map, "No such tensor allocated by the allocator."
        assert tensor.numel() >= numel, "Assert tensor buffer does is not large enough"
        self.id_to_params[tid].append((param, numel, tuple(shape)))
        param.data = tensor.narrow(0, 0, numel).view(shape)

    def release_tensor(self, tensor):
        self.release_tensor_with_id(id(tensor))

    def release_tensor_with_id(self, tensor_id):
        assert tensor_id in self.tensor_map, "Invalid tensor id"
        size = self.tensor_sizes.pop(tensor_id)
        del self.tensor_addresses[tensor_id], self.tensor_map[tensor_id]
        for param, _, _ in self.id_to_params.pop(tensor_id):
            param.data = torch.empty(0, dtype=param.dtype, device=param.device)
        self.total_free += size
        print_rank_0(f"Free before release {self.total_free - size}. Released {size}. Total free after {self.total_free}.")

    def print_allocation(self, resolution=200):
        cells = ["."] * resolution
        for addr, size, _ in self._live_sorted():
            for c in range(addr * resolution // self.size, min(resolution, -(-(addr + size) * resolution // self.size))):
                cells[c] = "x"
        print_rank_0("".join(cells))

    # ---- compaction
    def _defragment_memory(self):
        cursor = 0
        for addr, size, tid in self._live_sorted():
            if addr != cursor:
                src = self.buffer.narrow(0, addr, size)
                dst = self.buffer.narrow(0, cursor, size)
                dst.copy_(src.clone() if cursor + size > addr else src)  # ranges overlap when the slide is shorter than the block
                t = self.tensor_map[tid]
                t.data = dst
                for param, numel, shape in self.id_to_params[tid]:
                    param.data = dst.narrow(0, 0, numel).view(shape)
                self.tensor_addresses[tid] = cursor
            cursor += size
