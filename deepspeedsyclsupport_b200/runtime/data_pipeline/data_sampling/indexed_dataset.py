"""Memory-mapped indexed dataset: a `.bin` file of items stored back to back plus a `.idx` file that locates them.

On-disk format = Megatron-LM "MMIDIDX" version 1 (the format of the reference's data_sampling/indexed_dataset.py:369-617), so
files written by either implementation are interchangeable:

    .idx   9s magic "MMIDIDX\\0\\0" | u64 version = 1 | u8 dtype code | u64 n_items | u64 n_docs     (34-byte header)
           i32 sizes[n_items] | i64 byte_offsets[n_items] | i64 doc_idx[n_docs]
    .bin   item 0 | item 1 | ...

Reading maps both files once and returns zero-copy numpy views (the views keep the mapping alive; nothing is unmapped by
hand).  Writing streams items to the `.bin` file and keeps only their lengths; `finalize` writes the index.
"""
import os
import struct

import numpy as np
import torch

_HDR_MAGIC = b"MMIDIDX\x00\x00"
_HEADER = struct.Struct("<9sQBQQ")  # magic, version, dtype code, items, documents

# dtype codes of the format (6 and 7 both mean float64, as in the original table)
dtypes = {1: np.uint8, 2: np.int8, 3: np.int16, 4: np.int32, 5: np.int64, 6: np.float64, 7: np.double, 8: np.uint16, 9: np.uint32, 10: np.uint64}


def code(dtype):
    matches = [k for k, v in dtypes.items() if v == dtype]
    if not matches:
        raise ValueError(dtype)
    return matches[0]


def index_file_path(prefix_path):
    return prefix_path + ".idx"


def data_file_path(prefix_path):
    return prefix_path + ".bin"


def _missing(path):
    print(f"Dataset does not exist: {path}")


def infer_dataset_impl(path):
    if not MMapIndexedDataset.exists(path):
        return _missing(path)
    with open(index_file_path(path), "rb") as fd:
        return "mmap" if fd.read(len(_HDR_MAGIC)) == _HDR_MAGIC else None


def make_builder(out_file, impl="mmap", vocab_size=None):
    assert impl == "mmap", "only the mmap implementation is provided"
    small_vocab = vocab_size is not None and vocab_size < 65500
    return MMapIndexedDatasetBuilder(out_file, dtype=np.uint16 if small_vocab else np.int32)


def make_dataset(path, impl="mmap", skip_warmup=False):
    if not MMapIndexedDataset.exists(path):
        return _missing(path)
    chosen = infer_dataset_impl(path) if impl == "infer" else impl
    assert chosen == "mmap"
    return MMapIndexedDataset(path, skip_warmup)


def _map(path):
    """Read-only byte view of a file (empty files cannot be mapped)."""
    if os.path.getsize(path) == 0:
        return memoryview(b"")
    return memoryview(np.memmap(path, mode="r", order="C"))


class MMapIndexedDataset(torch.utils.data.Dataset):

    class Index(object):
        """The three arrays of an .idx file as views into its mapping."""

        def __init__(self, path, skip_warmup=False):
            with open(path, "rb") as fd:
                head = fd.read(_HEADER.size)
            magic, version, dtype_code, self._len, self._doc_count = _HEADER.unpack(head)
            assert magic == _HDR_MAGIC, "Index file doesn't match expected format. Make sure that --dataset-impl is configured properly."
            assert version == 1, f"unsupported index version {version}"
            self._dtype = dtypes[dtype_code]
            self._dtype_size = np.dtype(self._dtype).itemsize
            self._bin_buffer = view = _map(path)
            at = _HEADER.size
            self._sizes = np.frombuffer(view, dtype=np.int32, count=self._len, offset=at)
            at += self._sizes.nbytes
            self._pointers = np.frombuffer(view, dtype=np.int64, count=self._len, offset=at)
            at += self._pointers.nbytes
            self._doc_idx = np.frombuffer(view, dtype=np.int64, count=self._doc_count, offset=at)

        dtype = property(lambda self: self._dtype)
        sizes = property(lambda self: self._sizes)
        doc_idx = property(lambda self: self._doc_idx)

        def __getitem__(self, i):
            return self._pointers[i], self._sizes[i]

        def __len__(self):
            return self._len

    def __init__(self, path, skip_warmup=False):
        super().__init__()
        self._do_init(path, skip_warmup)

    def _do_init(self, path, skip_warmup):
        self._path = path
        self._index = self.Index(index_file_path(path), skip_warmup)
        self._bin_buffer = _map(data_file_path(path))

    # pickling (DataLoader workers): only the path travels, the worker maps the files itself
    def __getstate__(self):
        return self._path

    def __setstate__(self, state):
        self._do_init(state, skip_warmup=True)

    def __len__(self):
        return len(self._index)

    def _view(self, byte_offset, count):
        return np.frombuffer(self._bin_buffer, dtype=self._index.dtype, count=count, offset=byte_offset)

    def __getitem__(self, idx):
        if isinstance(idx, (int, np.integer)):
            where, count = self._index[int(idx)]
            return self._view(where, count)
        if not isinstance(idx, slice):
            raise TypeError(f"bad index {idx!r}")
        first, stop, step = idx.indices(len(self))
        if step != 1:
            raise ValueError("Slices into indexed_dataset must be contiguous")
        counts = self._index.sizes[first:stop]
        if len(counts) == 0:
            return []
        whole = self._view(self._index._pointers[first], int(counts.sum()))
        return np.split(whole, np.cumsum(counts)[:-1])

    def get(self, idx, offset=0, length=None):
        """Item `idx`, or `length` elements of it starting `offset` elements in."""
        where, count = self._index[idx]
        take = count - offset if length is None else length
        return self._view(where + offset * self._index._dtype_size, take)

    sizes = property(lambda self: self._index.sizes)
    doc_idx = property(lambda self: self._index.doc_idx)
    dtype = property(lambda self: self._index.dtype)
    supports_prefetch = property(lambda self: False)

    def size(self, index):
        return self._index.sizes[index]

    def get_doc_idx(self):
        return self._index._doc_idx

    def set_doc_idx(self, doc_idx_):
        self._index._doc_idx = doc_idx_

    @staticmethod
    def exists(path):
        return all(os.path.exists(f(path)) for f in (index_file_path, data_file_path))


class MMapIndexedDatasetBuilder(object):

    def __init__(self, out_file, dtype=np.int64):
        self._data_file = open(out_file, "wb")
        self._dtype = dtype
        self._sizes = []      # elements per item
        self._doc_idx = [0]   # item index at which each document starts (plus the end)

    def add_item_numpy(self, np_array):
        arr = np.ascontiguousarray(np_array, dtype=self._dtype)
        self._data_file.write(arr.tobytes())
        self._sizes.append(arr.size)

    def add_item(self, tensor):
        self.add_item_numpy(tensor.numpy() if torch.is_tensor(tensor) else np.asarray(tensor))

    def end_document(self):
        self._doc_idx.append(len(self._sizes))

    def merge_file_(self, another_file):
        """Append every item (and document boundary) of another dataset with the same dtype."""
        other = MMapIndexedDataset.Index(index_file_path(another_file))
        assert other.dtype == self._dtype, "datasets to merge must share a dtype"
        print(f"    concat {another_file} size={len(other.sizes)} for a total size of {len(other.sizes) + len(self._sizes)}")
        shift = len(self._sizes)
        self._sizes.extend(int(s) for s in other.sizes)
        self._doc_idx.extend(int(d) + shift for d in other.doc_idx[1:])
        with open(data_file_path(another_file), "rb") as src:
            for chunk in iter(lambda: src.read(1 << 24), b""):
                self._data_file.write(chunk)

    def finalize(self, index_file):
        self._data_file.close()
        sizes = np.asarray(self._sizes, dtype=np.int32)
        nbytes = sizes.astype(np.int64) * np.dtype(self._dtype).itemsize
        offsets = np.concatenate(([0], np.cumsum(nbytes)[:-1])).astype(np.int64) if len(sizes) else np.zeros(0, dtype=np.int64)
        with open(index_file, "wb") as fd:
            fd.write(_HEADER.pack(_HDR_MAGIC, 1, code(self._dtype), len(sizes), len(self._doc_idx)))
            for block in (sizes, offsets, np.asarray(self._doc_idx, dtype=np.int64)):
                fd.write(block.tobytes())
