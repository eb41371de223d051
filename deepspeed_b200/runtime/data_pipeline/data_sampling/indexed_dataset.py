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


# ---- legacy ("lazy" / "cached") format ------------------------------------------------------------------------
# fairseq / Megatron ``TNTIDX`` layout (reference :139-330): the index holds, for N items, the N+1 prefix offsets into
# a flat list of dimension sizes, the N+1 prefix element offsets into the data file, the flat sizes and the document
# boundaries. Items are read with positional reads (no shared file offset, so a dataset object can be used from several
# DataLoader threads), the whole index comes in with one read.
_LEGACY_MAGIC = b"TNTIDX\x00\x00"


def read_longs(f, n):
    a = np.empty(n, dtype=np.int64)
    f.readinto(a)
    return a


def write_longs(f, a):
    f.write(np.asarray(a, dtype=np.int64).tobytes())


def create_doc_idx(sizes):
    """Document boundaries for a size list in which an empty item ends a document."""
    return [0] + [i + 1 for i, s in enumerate(sizes) if s == 0]


class IndexedDataset(torch.utils.data.Dataset):
    """Reader of the legacy format: one positional read per item."""
    _HDR_MAGIC = _LEGACY_MAGIC

    def __init__(self, path):
        super().__init__()
        self.path = path
        self._fd = None
        self.read_index(path)

    def read_index(self, path):
        with open(index_file_path(path), "rb") as f:
            head = f.read(56)
            assert head[:8] == self._HDR_MAGIC, ("Index file doesn't match expected format. Make sure that "
                                                   "--dataset-impl is configured properly.")
            version, dcode, self.element_size, self._len, self.s, self.doc_count = struct.unpack("<6Q", head[8:])
            assert version == 1, f"unsupported index version {version}"
            self.dtype = dtypes[dcode][0]
            body = read_longs(f, 2 * (self._len + 1) + self.s + self.doc_count)
        n1 = self._len + 1
        self.dim_offsets, self.data_offsets = body[:n1], body[n1:2 * n1]
        self.sizes = body[2 * n1:2 * n1 + self.s]
        self.doc_idx = body[2 * n1 + self.s:]

    def read_data(self, path):
        self._fd = os.open(data_file_path(path), os.O_RDONLY)

    @property
    def data_file(self):
        return self._fd

    def check_index(self, i):
        if i < 0 or i >= self._len:
            raise IndexError("index out of range")

    def __del__(self):
        if getattr(self, "_fd", None) is not None:
            os.close(self._fd)
            self._fd = None

    def __getstate__(self):
        return self.path

    def __setstate__(self, path):
        self.__init__(path)

    def _read(self, first, last):
        """Elements of items ``[first, last)`` as one flat array."""
        if self._fd is None:
            self.read_data(self.path)
        lo, hi = int(self.data_offsets[first]), int(self.data_offsets[last])
        raw = os.pread(self._fd, (hi - lo) * self.element_size, lo * self.element_size)
        return np.frombuffer(raw, dtype=self.dtype).copy()

    def __getitem__(self, idx):
        if isinstance(idx, (int, np.integer)):
            i = int(idx)
            self.check_index(i)
            shape = self.sizes[self.dim_offsets[i]:self.dim_offsets[i + 1]]
            return self._read(i, i + 1).reshape(tuple(int(d) for d in shape))
        if isinstance(idx, slice):
            start, stop, step = idx.indices(len(self))
            if step != 1:
                raise ValueError("Slices into indexed_dataset must be contiguous")
            flat = self._read(start, stop)
            cuts = (self.data_offsets[start + 1:stop] - self.data_offsets[start]).astype(np.int64)
            return np.split(flat, cuts)
        raise TypeError(f"unsupported index {idx!r}")

    def __len__(self):
        return self._len

    def num_tokens(self, index):
        return self.sizes[index]

    def size(self, index):
        return self.sizes[index]

    @staticmethod
    def exists(path):
        return os.path.exists(index_file_path(path)) and os.path.exists(data_file_path(path))

    @property
    def supports_prefetch(self):
        return False


class IndexedCachedDataset(IndexedDataset):
    """Legacy reader that serves items from an in-memory cache filled by :meth:`prefetch`."""

    def __init__(self, path):
        super().__init__(path)
        self.cache = None
        self.cache_index = {}

    @property
    def supports_prefetch(self):
        return True

    def prefetch(self, indices):
        if all(i in self.cache_index for i in indices):
            return
        indices = sorted(set(int(i) for i in indices))
        lens = [int(self.data_offsets[i + 1] - self.data_offsets[i]) for i in indices]
        self.cache = np.empty(sum(lens), dtype=self.dtype)
        self.cache_index.clear()
        at = 0
        run_start = 0
        for k, i in enumerate(indices):  # neighbouring items come in with one read
            self.cache_index[i] = at
            at += lens[k]
            if k + 1 == len(indices) or indices[k + 1] != i + 1:
                first = indices[run_start]
                lo = self.cache_index[first]
                self.cache[lo:at] = self._read(first, i + 1)
                run_start = k + 1
        if self._fd is not None:
            os.close(self._fd)
            self._fd = None

    def __getitem__(self, idx):
        if isinstance(idx, (int, np.integer)):
            i = int(idx)
            self.check_index(i)
            shape = tuple(int(d) for d in self.sizes[self.dim_offsets[i]:self.dim_offsets[i + 1]])
            at = self.cache_index[i]
            n = int(self.data_offsets[i + 1] - self.data_offsets[i])
            return self.cache[at:at + n].reshape(shape).copy()
        if isinstance(idx, slice):
            return [self[i] for i in range(*idx.indices(len(self)))]
        raise TypeError(f"unsupported index {idx!r}")


class IndexedDatasetBuilder:
    """Writer of the legacy format."""

    def __init__(self, out_file, dtype=np.int32):
        self.out_file = open(out_file, "wb")
        self.dtype = dtype
        self.element_size = np.dtype(dtype).itemsize
        self.data_offsets, self.dim_offsets, self.sizes, self.doc_idx = [0], [0], [], [0]

    def add_item(self, tensor):
        arr = np.asarray(tensor.numpy() if torch.is_tensor(tensor) else tensor, dtype=self.dtype)
        self.out_file.write(arr.tobytes(order="C"))
        self.data_offsets.append(self.data_offsets[-1] + arr.size)
        self.sizes.extend(arr.shape)
        self.dim_offsets.append(self.dim_offsets[-1] + arr.ndim)

    def end_document(self):
        self.doc_idx.append(len(self.sizes))

    def merge_file_(self, another_file):
        other = IndexedDataset(another_file)
        assert np.dtype(other.dtype) == np.dtype(self.dtype)
        doc_offset = len(self.sizes)
        self.data_offsets.extend((self.data_offsets[-1] + other.data_offsets[1:]).tolist())
        self.dim_offsets.extend((self.dim_offsets[-1] + other.dim_offsets[1:]).tolist())
        self.sizes.extend(other.sizes.tolist())
        self.doc_idx.extend((doc_offset + other.doc_idx[1:]).tolist())
        with open(data_file_path(another_file), "rb") as f:
            shutil.copyfileobj(f, self.out_file, 1 << 22)

    def finalize(self, index_file):
        self.out_file.close()
        with open(index_file, "wb") as f:
            f.write(_LEGACY_MAGIC)
            f.write(struct.pack("<6Q", 1, code(self.dtype), self.element_size, len(self.data_offsets) - 1,
                                len(self.sizes), len(self.doc_idx)))
            for a in (self.dim_offsets, self.data_offsets, self.sizes, self.doc_idx):
                write_longs(f, a)


def get_available_dataset_impl():
    return ["lazy", "cached", "mmap"]


def make_builder(out_file, impl="mmap", vocab_size=None, dtype=None):
    if impl == "mmap":
        return MMapIndexedDatasetBuilder(out_file, dtype=dtype or best_fitting_dtype(vocab_size))
    return IndexedDatasetBuilder(out_file, **({"dtype": dtype} if dtype is not None else {}))


def make_dataset(path, impl="mmap", skip_warmup=False):
    if not IndexedDataset.exists(path):
        print(f"Dataset does not exist: {path}")
        print("Path should be a basename that both .idx and .bin can be appended to get full filenames.")
        return None
    if impl == "infer":
        impl = infer_dataset_impl(path)
    if impl == "lazy":
        return IndexedDataset(path)
    if impl == "cached":
        return IndexedCachedDataset(path)
    if impl == "mmap":
        return MMapIndexedDataset(path, skip_warmup)
    print(f"Unknown dataset implementation: {impl}")
    return None


def dataset_exists(path, impl="mmap"):
    return MMapIndexedDataset.exists(path) if impl == "mmap" else IndexedDataset.exists(path)


def infer_dataset_impl(path):
    if not IndexedDataset.exists(path):
        print(f"Dataset does not exist: {path}")
        return None
    with open(index_file_path(path), "rb") as f:
        magic = f.read(8)
    if magic == _LEGACY_MAGIC:
        return "cached"
    if magic == _MAGIC[:8]:
        return "mmap"
    return None
