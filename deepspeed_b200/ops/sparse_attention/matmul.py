"""Block-sparse matrix multiplication in its three forms (reference ``ops/sparse_attention/matmul.py::MatMul``).

``layout [H, M/block, N/block]`` marks the non-zero blocks of the sparse operand; a sparse tensor is stored as
``[B, nnz, block, block]`` with the blocks in row-major order of ``layout.nonzero()`` (head, block row, block column).

* ``sdd``: sparse = dense x dense - only the blocks the layout keeps are computed (``Q K^T`` restricted to the pattern)
* ``dsd``: dense = sparse x dense (``P V``)
* ``dds``: dense = dense x sparse

Each form is one gather of the operand blocks, one batched ``block x block`` GEMM over the ``nnz`` blocks (tensor
cores through cuBLAS batched GEMM) and, for the dense results, one ``index_add`` that sums a block row's partial
products. Autograd differentiates through these ops, so no hand-written backward modes are needed (the reference
derives them by swapping modes: d(sdd) = dsd / dds and so on)."""
import torch


class MatMul:

    def __init__(self, layout, block, mode, trans_a=False, trans_b=False, bench=False):
        if mode not in ("sdd", "dsd", "dds"):
            raise NotImplementedError("Supported modes are: sdd, dsd, dds")
        self.layout, self.block, self.mode = layout, block, mode
        self.trans_a, self.trans_b, self.bench = trans_a, trans_b, bench
        self.spdims = tuple(layout.shape)
        nz = layout.nonzero(as_tuple=False)  # [nnz, 3] rows (h, i, j), row-major
        self.nnz = nz.shape[0]
        self._nz = nz
        self._idx = {}

    def _index(self, device):
        key = str(device)
        if key not in self._idx:
            h, i, j = (self._nz[:, k].to(device) for k in range(3))
            self._idx[key] = (h, i, j)
        return self._idx[key]

    def _blocks(self, x, rows_block, cols_block):
        """Dense ``[B, H, R, C]`` -> ``[B, H, R/rb, C/cb, rb, cb]`` view of its blocks."""
        B, H, R, C = x.shape
        return x.reshape(B, H, R // rows_block, rows_block, C // cols_block, cols_block).permute(0, 1, 2, 4, 3, 5)

    def __call__(self, a, b):
        blk = self.block
        if self.mode == "sdd":
            if self.trans_a:
                a = a.transpose(-1, -2)
            if self.trans_b:
                b = b.transpose(-1, -2)
            h, i, j = self._index(a.device)
            B, H, M, K = a.shape
            a_rows = a.reshape(B, H, M // blk, blk, K)[:, h, i]  # [B, nnz, blk, K]
            b_cols = b.reshape(B, H, K, b.shape[-1] // blk, blk).permute(0, 1, 3, 2, 4)[:, h, j]  # [B, nnz, K, blk]
            return torch.matmul(a_rows, b_cols)
        if self.mode == "dsd":
            # a sparse [B, nnz, blk, blk] (block (i, j) of an [M, K] matrix), b dense [B, H, K, N]
            if self.trans_b:
                b = b.transpose(-1, -2)
            h, i, j = self._index(b.device)
            if self.trans_a:
                a, i, j = a.transpose(-1, -2), j, i
            B, H, K, N = b.shape
            b_rows = b.reshape(B, H, K // blk, blk, N)[:, h, j]  # [B, nnz, blk, N]
            part = torch.matmul(a, b_rows)  # [B, nnz, blk, N]
            n_rows = self.spdims[2] if self.trans_a else self.spdims[1]
            out = part.new_zeros(B, H * n_rows, blk, N)
            out.index_add_(1, h * n_rows + i, part)
            return out.view(B, H, n_rows * blk, N)
        # dds: a dense [B, H, M, K], b sparse (block (i, j) of a [K, N] matrix)
        if self.trans_a:
            a = a.transpose(-1, -2)
        h, i, j = self._index(a.device)
        if self.trans_b:
            b, i, j = b.transpose(-1, -2), j, i
        B, H, M, K = a.shape
        a_cols = a.reshape(B, H, M, K // blk, blk).permute(0, 1, 3, 2, 4)[:, h, i]  # [B, nnz, M, blk]
        part = torch.matmul(a_cols, b)  # [B, nnz, M, blk]
        n_cols = self.spdims[1] if self.trans_b else self.spdims[2]
        out = part.new_zeros(B, H * n_cols, M, blk)
        out.index_add_(1, h * n_cols + j, part)
        return out.view(B, H, n_cols, M, blk).permute(0, 1, 3, 2, 4).reshape(B, H, M, n_cols * blk)


def to_sparse(x, layout, block):
    """Dense ``[B, H, M, N]`` -> the ``[B, nnz, block, block]`` storage of its non-zero blocks."""
    nz = layout.nonzero(as_tuple=False).to(x.device)
    B, H, M, N = x.shape
    return x.reshape(B, H, M // block, block, N // block, block).permute(0, 1, 2, 4, 3, 5)[:, nz[:, 0], nz[:, 1],
                                                                                            nz[:, 2]]


def to_dense(x, layout, block, fill=0.0):
    """Inverse of :func:`to_sparse`; blocks outside the layout hold ``fill``."""
    nz = layout.nonzero(as_tuple=False).to(x.device)
    H, R, C = layout.shape
    B = x.shape[0]
    out = x.new_full((B, H, R, C, block, block), fill)
    out[:, nz[:, 0], nz[:, 1], nz[:, 2]] = x
    return out.permute(0, 1, 2, 4, 3, 5).reshape(B, H, R * block, C * block)
