"""Memory-mapped variable-length-row dataset in the Megatron ``MMIDIDX`` on-disk format, so files written by the
reference (``runtime/data_pipeline/data_sampling/indexed_dataset.py:369``) and by Megatron preprocessing load here
and vice versa.

``<prefix>.bin``: the rows back to back. ``<prefix>.idx``: magic ``MMIDIDX\\0\\0`` | u64 version=1 | u8 dtype code |
u64 n_rows | u64 n_docs | int32 sizes[n_rows] | int64 byte_offsets[n_rows] | int64 doc_idx[n_docs]."""
import os
import shutil
import struct

import numpy as np
import torch

_MAGIC = b"MMIDIDX\x00\x00"

dtypes = {
    1: (np.uint8, torch.uint8),
    2: (np.int8, torch.int8),
    3: (np.int16, torch.int16),
    4: (np.int32, torch.int32),
    5: (np.int64, torch.int64),
    6: (np.uint16, None),
    7: (np.uint32, None),
    8: (np.uint64, None),
}


def code(dtype):
    for c, (np_dt, torch_dt) in dtypes.items():
        if isinstance(dtype, torch.dtype):
            if dtype == torch_dt:
                return c
        elif np.dtype(dtype) == np.dtype(np_dt):
            return c
    raise ValueError(f"{dtype} not supported by the indexed dataset")


def index_file_path(prefix_path):
    return prefix_path + ".idx"


def data_file_path(prefix_path):
    return prefix_path + ".bin"


def best_fitting_dtype(vocab_size=None):
    return np.uint16 if vocab_size is not None and vocab_size < 65500 else np.int32


class _Index:

    def __init__(self, path):
        with open(path, "rb") as f:
            assert f.read(9) == _MAGIC, "index file does not match the MMIDIDX format"
            version, = struct.unpack("<Q", f.read(8))
            assert version == 1
            dcode, = struct.unpack("<B", f.read(1))
            self.dtype = dtypes[dcode][0]
            self.n_rows, self.n_docs = struct.unpack("<QQ", f.read(16))
            offset = f.tell()
        self._mmap = np.memmap(path, mode="r", order="C")
        buf = memoryview(self._mmap)
        self.sizes = np.frombuffer(buf, dtype=np.int32, count=self.n_rows, offset=offset)
        self.pointers = np.frombuffer(buf, dtype=np.int64, count=self.n_rows, offset=offset + self.sizes.nbytes)
        self.doc_idx = np.frombuffer(buf, dtype=np.int64, count=self.n_docs,
                                     offset=offset + self.sizes.nbytes + self.pointers.nbytes)

    @staticmethod
    def write(path, dtype, sizes, doc_idx):
        sizes = np.asarray(sizes, dtype=np.int64)
        pointers = np.zeros(len(sizes), dtype=np.int64)
        if len(sizes) > 1:
            np.cumsum(sizes[:-1] * np.dtype(dtype).itemsize, out=pointers[1:])
        with open(path, "wb") as f:
            f.write(_MAGIC)
            f.write(struct.pack("<Q", 1))
            f.write(struct.pack("<B", code(dtype)))
            f.write(struct.pack("<QQ", len(sizes), len(doc_idx)))
            f.write(sizes.astype(np.int32).tobytes(order="C"))
            f.write(pointers.tobytes(order="C"))
            f.write(np.asarray(doc_idx, dtype=np.int64).tobytes(order="C"))


class MMapIndexedDataset(torch.utils.data.Dataset):

    def __init__(self, path, skip_warmup=False):
        super().__init__()
        self._do_init(path, skip_warmup)

    def __getstate__(self):
        return self._path

    def __setstate__(self, state):
        self._do_init(state, skip_warmup=True)

    def _do_init(self, path, skip_warmup):
        self._path = path
        self._index = _Index(index_file_path(path))
        if not skip_warmup:  # touch the pages once so the first epoch is not bound by page faults
            with open(data_file_path(path), "rb") as f:
                while f.read(100 * 1024 * 1024):
                    pass
        if os.path.getsize(data_file_path(path)) == 0:
            self._bin_mmap, self._bin = None, memoryview(b"")
        else:
            self._bin_mmap = np.memmap(data_file_path(path), mode="r", order="C")
            self._bin = memoryview(self._bin_mmap)

    def __len__(self):
        return self._index.n_rows

    def __getitem__(self, idx):
        if isinstance(idx, (int, np.integer)):
            return np.frombuffer(self._bin, dtype=self._index.dtype, count=int(self._index.sizes[idx]),
                                 offset=int(self._index.pointers[idx]))
        if isinstance(idx, slice):
            start, stop, step = idx.indices(len(self))
            if step != 1:
                raise ValueError("Slices into indexed_dataset must be contiguous")
            sizes = self._index.sizes[idx]
            if len(sizes) == 0:
                return []
            flat = np.frombuffer(self._bin, dtype=self._index.dtype, count=int(sizes.sum()),
                                 offset=int(self._index.pointers[start]))
            return np.split(flat, np.cumsum(sizes)[:-1])
        raise TypeError(f"Unexpected type received for idx: {type(idx)}")

    def get(self, idx, offset=0, length=None):
        """A window of one row without materialising the whole row."""
        size = int(self._index.sizes[idx])
        length = size - offset if length is None else length
        ptr = int(self._index.pointers[idx]) + offset * np.dtype(self._index.dtype).itemsize
        return np.frombuffer(self._bin, dtype=self._index.dtype, count=length, offset=ptr)

    @property
    def sizes(self):
        return self._index.sizes

    def size(self, index):
        return self._index.sizes[index]

    @property
    def doc_idx(self):
        return self._index.doc_idx

    def get_doc_idx(self):
        return self._index.doc_idx

    def set_doc_idx(self, doc_idx_):
        self._index.doc_idx = doc_idx_

    @property
    def supports_prefetch(self):
        return False

    @property
    def dtype(self):
        return self._index.dtype

    @staticmethod
    def exists(path):
        return os.path.exists(index_file_path(path)) and os.path.exists(data_file_path(path))


class MMapIndexedDatasetBuilder:

    def __init__(self, out_file, dtype=np.int64):
        self._data_file = open(out_file, "wb")
        self._dtype = dtype
        self._sizes = []
        self._doc_idx = [0]

    def add_item(self, tensor):
        self.add_item_numpy(tensor.numpy() if isinstance(tensor, torch.Tensor) else np.asarray(tensor))

    def add_items(self, arr_list):
        for arr in arr_list:
            self.add_item(arr)

    def add_item_numpy(self, np_array):
        arr = np.ascontiguousarray(np_array, dtype=self._dtype)
        self._data_file.write(arr.tobytes(order="C"))
        self._sizes.append(arr.size)

    def end_document(self):
        self._doc_idx.append(len(self._sizes))

    def merge_file_(self, another_file):
        other = _Index(index_file_path(another_file))
        assert other.dtype == self._dtype
        base = len(self._sizes)
        self._sizes.extend(int(s) for s in other.sizes)
        self._doc_idx.extend(int(d) + base for d in other.doc_idx[1:])
        with open(data_file_path(another_file), "rb") as f:
            shutil.copyfileobj(f, self._data_file)

    def finalize(self, index_file):
        self._data_file.close()
        _Index.write(index_file, self._dtype, self._sizes, self._doc_idx)


def make_builder(out_file, impl="mmap", vocab_size=None, dtype=None):
    return MMapIndexedDatasetBuilder(out_file, dtype=dtype or best_fitting_dtype(vocab_size))


def make_dataset(path, impl="mmap", skip_warmup=False):
    if not MMapIndexedDataset.exists(path):
        print(f"Dataset does not exist: {path}")
        return None
    return MMapIndexedDataset(path, skip_warmup)


def dataset_exists(path, impl="mmap"):
    return MMapIndexedDataset.exists(path)


def infer_dataset_impl(path):
    if MMapIndexedDataset.exists(path):
        with open(index_file_path(path), "rb") as f:
            if f.read(9)[:8] == _MAGIC[:8]:
                return "mmap"
    return None
