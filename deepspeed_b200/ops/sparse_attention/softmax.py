"""Softmax over the rows of a block-sparse matrix (reference ``ops/sparse_attention/softmax.py::Softmax``).

``x [B, nnz, block, block]`` in the storage of :mod:`.matmul`. A row of the logical matrix is spread over the blocks of
its block row; the normalisation runs over all of them: row maxima and sums are reduced per (head, block row) with
``scatter_reduce`` / ``index_add`` keyed by the block's row id, so no dense ``[L, L]`` matrix is formed. ``scale``,
relative position embedding, key padding mask and attention mask are applied before the normalisation, with the
reference's ``add`` / ``mul`` mask modes."""
import torch


class Softmax:

    def __init__(self, layout, block, bench=False):
        self.layout, self.block, self.bench = layout, block, bench
        self.spdims = tuple(layout.shape)
        self._nz = layout.nonzero(as_tuple=False)
        self._idx = {}

    def _index(self, device):
        key = str(device)
        if key not in self._idx:
            self._idx[key] = tuple(self._nz[:, k].to(device) for k in range(3))
        return self._idx[key]

    @staticmethod
    def _additive(mask, mode, dtype):
        mask = mask.to(dtype)
        return mask if mode == "add" else (1.0 - mask) * torch.finfo(dtype).min

    def __call__(self, x, scale=1.0, rpe=None, key_padding_mask=None, attn_mask=None, key_padding_mask_mode="add",
                 attn_mask_mode="add"):
        blk = self.block
        H, nbr, nbc = self.spdims
        h, i, j = self._index(x.device)
        B, nnz = x.shape[0], x.shape[1]
        s = x.float() * scale
        cols = (j.view(-1, 1) * blk + torch.arange(blk, device=x.device)).view(nnz, 1, blk)  # key position per column
        rows = (i.view(-1, 1) * blk + torch.arange(blk, device=x.device)).view(nnz, blk, 1)
        if rpe is not None:
            r = rpe if rpe.dim() == 3 else rpe.unsqueeze(0).expand(H, -1, -1)
            s = s + r[h.view(-1, 1, 1), rows, cols].float().unsqueeze(0)
        if attn_mask is not None:
            am = self._additive(attn_mask, attn_mask_mode, torch.float32)
            s = s + am[rows, cols].unsqueeze(0)
        if key_padding_mask is not None:
            kp = self._additive(key_padding_mask, key_padding_mask_mode, torch.float32)  # [B, L]
            s = s + kp[:, cols.view(nnz, blk)].view(B, nnz, 1, blk)
        seg = h * nbr + i  # which (head, block row) a block belongs to
        n_seg = H * nbr
        blk_max = s.amax(dim=-1)  # [B, nnz, blk]
        row_max = torch.full((B, n_seg, blk), float("-inf"), device=x.device, dtype=torch.float32)
        row_max = row_max.scatter_reduce(1, seg.view(1, -1, 1).expand(B, nnz, blk), blk_max, reduce="amax",
                                         include_self=True)
        e = torch.exp(s - row_max[:, seg].unsqueeze(-1))
        row_sum = torch.zeros(B, n_seg, blk, device=x.device, dtype=torch.float32)
        row_sum.index_add_(1, seg, e.sum(dim=-1))
        return (e / row_sum[:, seg].unsqueeze(-1)).to(x.dtype)
