"""Memory-mapped indexed dataset (.bin + .idx) in the Megatron-LM "MMIDIDX" on-disk format, so index files produced by
either implementation are interchangeable (reference: data_sampling/indexed_dataset.py:369-617).

    .idx : b"MMIDIDX\\0\\0" | u64 version=1 | u8 dtype-code | u64 n_items | u64 n_docs |
           i32 sizes[n_items] | i64 byte_offsets[n_items] | i64 doc_idx[n_docs]
    .bin : the items back to back

The reader keeps one np.memmap per file and hands out zero-copy views; the builder streams items to disk and keeps only
the size list in memory."""
import os
import struct
from functools import lru_cache

import numpy as np
import torch

_HDR_MAGIC = b"MMIDIDX\x00\x00"

dtypes = {1: np.uint8, 2: np.int8, 3: np.int16, 4: np.int32, 5: np.int64, 6: np.float64, 7: np.double, 8: np.uint16, 9: np.uint32, 10: np.uint64}


def code(dtype):
    for k, v in dtypes.items():
        if v == dtype:
            return k
    raise ValueError(dtype)


def index_file_path(prefix_path):
    return prefix_path + ".idx"


def data_file_path(prefix_path):
    return prefix_path + ".bin"


def infer_dataset_impl(path):
    if MMapIndexedDataset.exists(path):
        with open(index_file_path(path), "rb") as f:
            return "mmap" if f.read(9) == _HDR_MAGIC else None
    print(f"Dataset does not exist: {path}")
    return None


def make_builder(out_file, impl="mmap", vocab_size=None):
    assert impl == "mmap", "only the mmap implementation is provided"
    return MMapIndexedDatasetBuilder(out_file, dtype=(np.uint16 if vocab_size is not None and vocab_size < 65500 else np.int32))


def make_dataset(path, impl="mmap", skip_warmup=False):
    if not MMapIndexedDataset.exists(path):
        print(f"Dataset does not exist: {path}")
        return None
    if impl == "infer":
        impl = infer_dataset_impl(path)
    assert impl == "mmap"
    return MMapIndexedDataset(path, skip_warmup)


class MMapIndexedDataset(torch.utils.data.Dataset):

    class Index(object):

        def __init__(self, path, skip_warmup=False):
            with open(path, "rb") as stream:
                assert stream.read(9) == _HDR_MAGIC, ("Index file doesn't match expected format. Make sure that --dataset-impl is configured properly.")
                version, = struct.unpack("<Q", stream.read(8))
                assert version == 1
                dtype_code, = struct.unpack("<B", stream.read(1))
                self._dtype = dtypes[dtype_code]
                self._dtype_size = self._dtype().itemsize
                self._len, = struct.unpack("<Q", stream.read(8))
                self._doc_count, = struct.unpack("<Q", stream.read(8))
                offset = stream.tell()
            self._bin_buffer_mmap = np.memmap(path, mode="r", order="C")
            self._bin_buffer = memoryview(self._bin_buffer_mmap)
            self._sizes = np.frombuffer(self._bin_buffer, dtype=np.int32, count=self._len, offset=offset)
            self._pointers = np.frombuffer(self._bin_buffer, dtype=np.int64, count=self._len, offset=offset + self._sizes.nbytes)
            self._doc_idx = np.frombuffer(self._bin_buffer, dtype=np.int64, count=self._doc_count,
                                          offset=offset + self._sizes.nbytes + self._pointers.nbytes)

        @property
        def dtype(self):
            return self._dtype

        @property
        def sizes(self):
            return self._sizes

        @property
        def doc_idx(self):
            return self._doc_idx

        @lru_cache(maxsize=8)
        def __getitem__(self, i):
            return self._pointers[i], self._sizes[i]

        def __len__(self):
            return self._len

    def __init__(self, path, skip_warmup=False):
        super().__init__()
        self._path = None
        self._index = None
        self._bin_buffer = None
        self._do_init(path, skip_warmup)

    def __getstate__(self):
        return self._path

    def __setstate__(self, state):
        self._do_init(state, skip_warmup=True)

    def _do_init(self, path, skip_warmup):
        self._path = path
        self._index = self.Index(index_file_path(self._path), skip_warmup)
        if os.path.getsize(data_file_path(self._path)) == 0:
            self._bin_buffer_mmap = None
            self._bin_buffer = memoryview(b"")
        else:
            self._bin_buffer_mmap = np.memmap(data_file_path(self._path), mode="r", order="C")
            self._bin_buffer = memoryview(self._bin_buffer_mmap)

    # no explicit munmap: rows handed out are zero-copy views into the mapping, which must stay valid for as long as any
    # of them is alive (numpy keeps the mapping referenced through the view's base)

    def __len__(self):
        return len(self._index)

    def __getitem__(self, idx):
        if isinstance(idx, (int, np.integer)):
            ptr, size = self._index[int(idx)]
            return np.frombuffer(self._bin_buffer, dtype=self._index.dtype, count=size, offset=ptr)
        if isinstance(idx, slice):
            start, stop, step = idx.indices(len(self))
            if step != 1:
                raise ValueError("Slices into indexed_dataset must be contiguous")
            ptr = self._index._pointers[start]
            sizes = self._index._sizes[idx]
            offsets = list(np.cumsum(sizes))
            total_size = sum(sizes)
            np_array = np.frombuffer(self._bin_buffer, dtype=self._index.dtype, count=total_size, offset=ptr)
            return np.split(np_array, offsets[:-1])
        raise TypeError(f"bad index {idx!r}")

    def get(self, idx, offset=0, length=None):
        """Retrieves a single item from the dataset with the option to only return a portion of the item."""
        ptr, size = self._index[idx]
        if length is None:
            length = size - offset
        ptr += offset * np.dtype(self._index.dtype).itemsize
        return np.frombuffer(self._bin_buffer, dtype=self._index.dtype, count=length, offset=ptr)

    @property
    def sizes(self):
        return self._index.sizes

    def size(self, index):
        return self._index.sizes[index]

    @property
    def doc_idx(self):
        return self._index.doc_idx

    def get_doc_idx(self):
        return self._index._doc_idx

    def set_doc_idx(self, doc_idx_):
        self._index._doc_idx = doc_idx_

    @property
    def supports_prefetch(self):
        return False

    @property
    def dtype(self):
        return self._index.dtype

    @staticmethod
    def exists(path):
        return os.path.exists(index_file_path(path)) and os.path.exists(data_file_path(path))


class MMapIndexedDatasetBuilder(object):

    def __init__(self, out_file, dtype=np.int64):
        self._data_file = open(out_file, "wb")
        self._dtype = dtype
        self._sizes = []
        self._doc_idx = [0]

    def add_item(self, tensor):
        np_array = np.array(tensor.numpy() if torch.is_tensor(tensor) else tensor, dtype=self._dtype)
        self._data_file.write(np_array.tobytes(order="C"))
        self._sizes.append(np_array.size)

    def add_item_numpy(self, np_array):
        if np_array.dtype != self._dtype:
            np_array = np_array.astype(self._dtype)
        self._data_file.write(np_array.tobytes(order="C"))
        self._sizes.append(np_array.size)

    def end_document(self):
        self._doc_idx.append(len(self._sizes))

    def merge_file_(self, another_file):
        index = MMapIndexedDataset.Index(index_file_path(another_file))
        assert index.dtype == self._dtype
        total_len = len(index.sizes) + len(self._sizes)
        print(f"    concat {another_file} size={len(index.sizes)} for a total size of {total_len}")
        offset = len(self._sizes)
        self._sizes.extend(index.sizes)
        self._doc_idx.extend((offset + index.doc_idx)[1:])
        with open(data_file_path(another_file), "rb") as f:
            while True:
                chunk = f.read(1 << 24)
                if not chunk:
                    break
                self._data_file.write(chunk)

    def finalize(self, index_file):
        self._data_file.close()
        sizes = np.array(self._sizes, dtype=np.int32)
        itemsize = np.dtype(self._dtype).itemsize
        pointers = np.zeros(len(sizes), dtype=np.int64)
        if len(sizes) > 1:
            np.cumsum(sizes[:-1].astype(np.int64) * itemsize, out=pointers[1:])
        with open(index_file, "wb") as f:
            f.write(_HDR_MAGIC)
            f.write(struct.pack("<Q", 1))
            f.write(struct.pack("<B", code(self._dtype)))
            f.write(struct.pack("<Q", len(sizes)))
            f.write(struct.pack("<Q", len(self._doc_idx)))
            f.write(sizes.tobytes(order="C"))
            f.write(pointers.tobytes(order="C"))
            f.write(np.array(self._doc_idx, dtype=np.int64).tobytes(order="C"))
