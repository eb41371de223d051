"""Row-sparse gradient container for embedding layers + sparse all-reduce (reference ``runtime/sparse_tensor.py:13``,
engine ``sparse_allreduce`` :3815: all-gather indices and values instead of all-reducing the dense tensor)."""
import torch

from .. import comm as dist


class SparseTensor:

    def __init__(self, dense_tensor=None):
        self.orig_dense_tensor = dense_tensor
        self.dtype = dense_tensor.dtype if dense_tensor is not None else None
        if dense_tensor is not None:
            if dense_tensor.is_sparse:
                dense_tensor = dense_tensor.coalesce()
                self.indices = dense_tensor.indices().flatten()
                self.values = dense_tensor.values()
            else:
                result = torch.sum(dense_tensor, dim=1)
                self.indices = result.nonzero().flatten()
                self.values = dense_tensor[self.indices]
            self.dense_size = list(dense_tensor.size())
        else:
            self.indices = self.values = self.dense_size = None

    def to_coo_tensor(self):
        return torch.sparse_coo_tensor(self.indices.unsqueeze(0), self.values, self.dense_size)

    @staticmethod
    def type():
        return "deepspeed.SparseTensor"

    def to_dense(self):
        it = self.indices.unsqueeze(1)
        full_indices = torch.cat([it for _ in range(self.dense_size[1])], dim=1)
        return self.values.new_zeros(self.dense_size).scatter_add_(0, full_indices, self.values)

    def sparse_size(self):
        index_size = list(self.indices.size())
        value_size = list(self.values.size())
        dense_size = self.dense_size[0] * self.dense_size[1]
        return index_size[0] + value_size[0] * value_size[1], dense_size

    def add(self, b):
        assert self.dense_size == b.dense_size
        self.indices = torch.cat([self.indices, b.indices])
        self.values = torch.cat([self.values, b.values])

    def __str__(self):
        sparse_size, dense_size = self.sparse_size()
        return (f"DeepSpeed.SparseTensor(indices_size={self.indices.size()}, values_size={self.values.size()}, "
                f"dense_size={self.dense_size}, device={self.indices.device}, "
                f"reduction_factor={dense_size / sparse_size})")

    __repr__ = __str__


def sparse_allreduce(sparse, group=None, average=True):
    """All-gather (padded) indices/values of every rank; the result holds the concatenation (duplicates summed by
    ``to_dense``)."""
    world = dist.get_world_size(group)
    if world == 1:
        return sparse
    values = sparse.values / world if average else sparse.values
    n = torch.tensor([sparse.indices.numel()], device=sparse.indices.device)
    sizes = [torch.zeros_like(n) for _ in range(world)]
    dist.all_gather(sizes, n, group=group)
    mx = int(max(int(s) for s in sizes))
    pad_i = sparse.indices.new_zeros(mx)
    pad_i[:sparse.indices.numel()] = sparse.indices
    pad_v = values.new_zeros((mx, ) + tuple(values.shape[1:]))
    pad_v[:values.shape[0]] = values
    gi = [torch.empty_like(pad_i) for _ in range(world)]
    gv = [torch.empty_like(pad_v) for _ in range(world)]
    dist.all_gather(gi, pad_i, group=group)
    dist.all_gather(gv, pad_v, group=group)
    sparse.indices = torch.cat([g[:int(s)] for g, s in zip(gi, sizes)])
    sparse.values = torch.cat([g[:int(s)] for g, s in zip(gv, sizes)])
    return sparse
