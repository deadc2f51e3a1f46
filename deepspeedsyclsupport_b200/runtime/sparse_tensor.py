"""SparseTensor: row-sparse (index, values) form of an embedding gradient for the `sparse_gradients` all-reduce path
(parity: deepspeed/runtime/sparse_tensor.py).  Only the rows that received gradient are communicated."""
import torch


class SparseTensor(object):
    """Compressed Sparse Tensor"""

    def __init__(self, dense_tensor=None):
        self.orig_dense_tensor = dense_tensor
        self.dtype = self.orig_dense_tensor.dtype if dense_tensor is not None else None
        if dense_tensor is None:
            self.indices = self.values = self.dense_size = None
            return
        if dense_tensor.is_sparse:
            dense_tensor = dense_tensor.coalesce()
            self.indices = dense_tensor.indices().flatten()
            self.values = dense_tensor.values()
        else:
            rows = dense_tensor.reshape(dense_tensor.shape[0], -1).abs().sum(dim=1)
            self.indices = rows.nonzero(as_tuple=False).flatten()
            self.values = dense_tensor[self.indices]
        self.dense_size = list(dense_tensor.size())

    def to_coo_tensor(self):
        return torch.sparse_coo_tensor(self.indices.unsqueeze(0), self.values, self.dense_size)

    @staticmethod
    def type():
        return "deepspeed.SparseTensor"

    def to_dense(self):
        out = torch.zeros(self.dense_size, dtype=self.values.dtype, device=self.values.device)
        return out.index_add_(0, self.indices, self.values)

    def sparse_size(self):
        return self.indices.numel() + self.values.numel(), int(torch.tensor(self.dense_size).prod())

    def add(self, b):
        assert self.dense_size == b.dense_size
        self.indices = torch.cat([self.indices, b.indices])
        self.values = torch.cat([self.values, b.values])

    def __str__(self):
        sparse_size, dense_size = self.sparse_size()
        return "DeepSpeed.SparseTensor(indices_size={}, values_size={}, dense_size={}, device={}, reduction_factor={})".format(
            self.indices.size(), self.values.size(), self.dense_size, self.indices.device, dense_size / max(sparse_size, 1))

    def __repr__(self):
        return self.__str__()


def sparse_allreduce(sparse: SparseTensor, group=None):
    """All-reduce (average) of a row-sparse gradient: ranks exchange their (padded) index / value lists with
    all_gather and concatenate -- duplicates are summed when the result is densified (engine.py:2447-2516 upstream)."""
    from .. import comm as dist
    world = dist.get_world_size(group)
    n = torch.tensor([sparse.indices.numel()], device=sparse.indices.device)
    sizes = [torch.zeros_like(n) for _ in range(world)]
    dist.all_gather(sizes, n, group=group)
    max_n = int(max(s.item() for s in sizes))
    pad = max_n - sparse.indices.numel()
    idx = torch.cat([sparse.indices, sparse.indices.new_zeros(pad)])
    val = torch.cat([sparse.values, sparse.values.new_zeros(pad, *sparse.values.shape[1:])])
    all_idx = [torch.empty_like(idx) for _ in range(world)]
    all_val = [torch.empty_like(val) for _ in range(world)]
    dist.all_gather(all_idx, idx, group=group)
    dist.all_gather(all_val, val, group=group)
    sparse.indices = torch.cat([i[:int(s.item())] for i, s in zip(all_idx, sizes)])
    sparse.values = torch.cat([v[:int(s.item())] for v, s in zip(all_val, sizes)]) / world
    return sparse
