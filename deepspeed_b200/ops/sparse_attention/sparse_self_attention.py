"""Block-sparse self attention (reference ``ops/sparse_attention/sparse_self_attention.py:12`` on Triton block-sparse
matmul/softmax kernels, ``matmul.py`` / ``softmax.py``).

Computation here is *gather based*: for every (head, query block) the active key/value blocks listed by the layout are
gathered into a dense [block, K_max * block] problem, so FLOPs and memory scale with the number of active blocks, and
the three steps (QK^T, softmax, PV) run as batched tensor-core matmuls. Padding slots, key-padding masks, additive
attention masks and relative position embeddings are folded in before the softmax (same semantics as the reference's
``Softmax`` op with ``rpe`` / ``key_padding_mask`` / ``attn_mask``)."""
import torch
import torch.nn as nn

from .sparsity_config import SparsityConfig


class _LayoutPlan:
    """Padded per-row lists of active key blocks of a layout."""

    def __init__(self, layout, device):
        H, nb, _ = layout.shape
        counts = layout.sum(-1)
        self.kmax = int(counts.max())
        order = torch.argsort(layout, dim=-1, descending=True, stable=True)[..., :self.kmax]  # active columns first
        valid = torch.arange(self.kmax).view(1, 1, -1) < counts.unsqueeze(-1)
        self.idx = torch.where(valid, order, torch.zeros_like(order)).to(device)
        self.valid = valid.to(device)
        self.nb = nb


class SparseSelfAttention(nn.Module):
    ops = dict()

    def __init__(self, sparsity_config=SparsityConfig(num_heads=4), key_padding_mask_mode="add", attn_mask_mode="mul",
                 max_seq_length=2048):
        super().__init__()
        self.sparsity_config = sparsity_config
        master = sparsity_config.make_layout(max_seq_length)
        self.register_buffer("master_layout", master)
        self._need_layout_synchronization = True
        self.key_padding_mask_mode, self.attn_mask_mode = key_padding_mask_mode, attn_mask_mode
        self._plans = {}

    def get_layout(self, L):
        if self._need_layout_synchronization and torch.distributed.is_available() and \
                torch.distributed.is_initialized():
            torch.distributed.broadcast(self.master_layout, src=0)  # random layouts must agree across ranks
            self._need_layout_synchronization = False
        if L % self.sparsity_config.block != 0:
            raise ValueError(f"Sequence Length, {L}, needs to be dividable by Block size {self.sparsity_config.block}!")
        nb = L // self.sparsity_config.block
        return self.master_layout[..., :nb, :nb].cpu()

    def _plan(self, L, device):
        key = (L, str(device))
        if key not in self._plans:
            self._plans[key] = _LayoutPlan(self.get_layout(L), device)
        return self._plans[key]

    @staticmethod
    def _as_additive(mask, mode, dtype):
        if mask is None:
            return None
        mask = mask.to(dtype)
        return mask if mode == "add" else (1.0 - mask) * torch.finfo(dtype).min

    def forward(self, query, key, value, rpe=None, key_padding_mask=None, attn_mask=None):
        """query/key/value: [B, H, L, D]. rpe: [L, L] or [H, L, L]; key_padding_mask: [B, L]; attn_mask: [L, L]."""
        assert query.dtype in (torch.half, torch.bfloat16, torch.float32), "unsupported dtype"
        B, H, L, D = query.shape
        if query.shape != key.shape or key.shape != value.shape:
            raise NotImplementedError("only self-attention is supported for now")
        blk = self.sparsity_config.block
        plan = self._plan(L, query.device)
        nb, km = plan.nb, plan.kmax
        idx = plan.idx                                            # [H, nb, km]
        scale = float(D)**-0.5
        qb = query.reshape(B, H, nb, blk, D)
        kb = key.reshape(B, H, nb, blk, D)
        vb = value.reshape(B, H, nb, blk, D)
        gi = idx.view(1, H, nb * km, 1, 1).expand(B, H, nb * km, blk, D)
        kg = torch.gather(kb, 2, gi).view(B, H, nb, km * blk, D)
        vg = torch.gather(vb, 2, gi).view(B, H, nb, km * blk, D)
        scores = torch.matmul(qb, kg.transpose(-1, -2)).float() * scale   # [B, H, nb, blk, km*blk]
        # absolute key position of every gathered column / query position of every row
        kpos = (idx.unsqueeze(-1) * blk + torch.arange(blk, device=idx.device)).view(H, nb, km * blk)
        neg = torch.finfo(torch.float32).min
        scores = scores.masked_fill(~plan.valid.repeat_interleave(blk, dim=-1).view(1, H, nb, 1, km * blk), neg)
        if rpe is not None:
            r = rpe if rpe.dim() == 3 else rpe.unsqueeze(0).expand(H, L, L)
            rb = r.reshape(H, nb, blk, L)
            scores = scores + torch.gather(rb, 3, kpos.unsqueeze(2).expand(H, nb, blk, km * blk)).unsqueeze(0).float()
        if attn_mask is not None:
            am = self._as_additive(attn_mask, self.attn_mask_mode, torch.float32).reshape(nb, blk, L)
            scores = scores + torch.gather(am.unsqueeze(0).expand(H, nb, blk, L), 3,
                                           kpos.unsqueeze(2).expand(H, nb, blk, km * blk)).unsqueeze(0)
        if key_padding_mask is not None:
            kp = self._as_additive(key_padding_mask, self.key_padding_mask_mode, torch.float32)  # [B, L]
            scores = scores + kp[:, kpos.reshape(-1)].view(B, H, nb, 1, km * blk)
        probs = torch.softmax(scores, dim=-1).to(query.dtype)
        out = torch.matmul(probs, vg)                              # [B, H, nb, blk, D]
        return out.reshape(B, H, L, D)
