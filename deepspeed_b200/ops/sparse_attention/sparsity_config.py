"""Block-sparse attention layouts (reference ``ops/sparse_attention/sparsity_config.py``): Dense, Fixed (Sparse
Transformers), Variable, BigBird, Block-Sparse Longformer, Local Sliding Window. A layout is a 0/1 tensor
``[num_heads, num_blocks, num_blocks]``; the builders are tensor expressions over block-index grids instead of
per-element loops."""
import random

import torch


def _grid(nb):
    r = torch.arange(nb).view(nb, 1)
    c = torch.arange(nb).view(1, nb)
    return r, c


class SparsityConfig:

    def __init__(self, num_heads, block=16, different_layout_per_head=False):
        self.num_heads = num_heads
        self.block = block
        self.different_layout_per_head = different_layout_per_head
        self.num_layout_heads = num_heads if different_layout_per_head else 1

    def setup_layout(self, seq_len):
        if seq_len % self.block != 0:
            raise ValueError(f"Sequence Length, {seq_len}, needs to be dividable by Block size {self.block}!")
        nb = seq_len // self.block
        return torch.zeros((self.num_heads, nb, nb), dtype=torch.int64)

    def check_and_propagate_first_head_layout(self, layout):
        if not self.different_layout_per_head:
            layout[1:self.num_heads] = layout[0]
        return layout

    def _sliding_window(self, nb, num_window_blocks):
        if nb < num_window_blocks:
            raise ValueError(f"Number of sliding window blocks, {num_window_blocks}, must be smaller than overall "
                             f"number of blocks in a row, {nb}!")
        r, c = _grid(nb)
        return ((c - r).abs() <= num_window_blocks // 2).long()

    def _random(self, nb, n_random, causal):
        if nb < n_random:
            raise ValueError(f"Number of random blocks, {n_random}, must be smaller than overall number of blocks in "
                             f"a row, {nb}!")
        out = torch.zeros(nb, nb, dtype=torch.int64)
        for row in range(nb):
            pool = range(0, row + 1) if causal else range(0, nb)
            out[row, random.sample(pool, min(n_random, len(pool)))] = 1
        return out


class DenseSparsityConfig(SparsityConfig):

    def make_layout(self, seq_len):
        layout = self.setup_layout(seq_len)
        layout[:] = 1
        return layout


class FixedSparsityConfig(SparsityConfig):

    def __init__(self, num_heads, block=16, different_layout_per_head=False, num_local_blocks=4, num_global_blocks=1,
                 attention="bidirectional", horizontal_global_attention=False, num_different_global_patterns=1):
        super().__init__(num_heads, block, different_layout_per_head)
        if num_local_blocks % num_global_blocks != 0:
            raise ValueError(f"Number of blocks in a local window, {num_local_blocks}, must be dividable by number of "
                             f"global blocks, {num_global_blocks}!")
        if attention not in ("unidirectional", "bidirectional"):
            raise NotImplementedError('only "uni/bi-directional" attentions are supported for now!')
        if attention != "bidirectional" and horizontal_global_attention:
            raise ValueError('only "bi-directional" attentions can support horizontal global attention!')
        if num_different_global_patterns > 1 and not different_layout_per_head:
            raise ValueError("Number of different layouts cannot be more than one when you have set a single layout "
                             "for all heads! Set different_layout_per_head to True.")
        if num_different_global_patterns > num_local_blocks // num_global_blocks:
            raise ValueError(f"Number of layout versions (num_different_global_patterns), "
                             f"{num_different_global_patterns}, cannot be larger than number of local window blocks "
                             f"divided by number of global blocks, {num_local_blocks // num_global_blocks}!")
        self.num_local_blocks, self.num_global_blocks = num_local_blocks, num_global_blocks
        self.attention, self.horizontal_global_attention = attention, horizontal_global_attention
        self.num_different_global_patterns = num_different_global_patterns

    def set_local_layout(self, h, layout):
        nb = layout.shape[1]
        r, c = _grid(nb)
        same_window = (r // self.num_local_blocks) == (c // self.num_local_blocks)
        if self.attention == "unidirectional":
            same_window = same_window & (c <= r)
        layout[h] |= same_window.long()
        return layout

    def set_global_layout(self, h, layout):
        nb = layout.shape[1]
        L, G = self.num_local_blocks, self.num_global_blocks
        first = L - (1 + h % self.num_different_global_patterns) * G
        end = nb - (nb % L)
        starts = list(range(first, end, L))
        if end < nb:
            starts.append(min(end + first, nb - G))
        for i in starts:
            first_row = 0 if self.attention == "bidirectional" else i
            layout[h, first_row:, i:i + G] = 1
            if self.horizontal_global_attention:
                layout[h, i:i + G, :] = 1
        return layout

    def make_layout(self, seq_len):
        layout = self.setup_layout(seq_len)
        for h in range(self.num_layout_heads):
            layout = self.set_global_layout(h, self.set_local_layout(h, layout))
        return self.check_and_propagate_first_head_layout(layout)


class VariableSparsityConfig(SparsityConfig):

    def __init__(self, num_heads, block=16, different_layout_per_head=False, num_random_blocks=0,
                 local_window_blocks=(4, ), global_block_indices=(0, ), global_block_end_indices=None,
                 attention="bidirectional", horizontal_global_attention=False):
        super().__init__(num_heads, block, different_layout_per_head)
        if global_block_end_indices is not None:
            if len(global_block_indices) != len(global_block_end_indices):
                raise ValueError("Global block start and end index lists must have the same length!")
            for s, e in zip(global_block_indices, global_block_end_indices):
                if s >= e:
                    raise ValueError(f"Global block start index, {s}, must be smaller than global block end index, {e}!")
        if attention not in ("unidirectional", "bidirectional"):
            raise NotImplementedError('only "uni/bi-directional" attentions are supported for now!')
        if attention != "bidirectional" and horizontal_global_attention:
            raise ValueError('only "bi-directional" attentions can support horizontal global attention!')
        self.num_random_blocks = num_random_blocks
        self.local_window_blocks = list(local_window_blocks)
        self.global_block_indices = list(global_block_indices)
        self.global_block_end_indices = list(global_block_end_indices) if global_block_end_indices is not None else None
        self.attention, self.horizontal_global_attention = attention, horizontal_global_attention

    def set_random_layout(self, h, layout):
        layout[h] |= self._random(layout.shape[1], self.num_random_blocks, causal=False)
        return layout

    def set_local_layout(self, h, layout):
        nb = layout.shape[1]
        window_of = torch.empty(nb, dtype=torch.int64)
        start, wid = 0, 0
        sizes = list(self.local_window_blocks)
        while start < nb:
            size = sizes[min(wid, len(sizes) - 1)]
            window_of[start:start + size] = wid
            start += size
            wid += 1
        same = window_of.view(nb, 1) == window_of.view(1, nb)
        if self.attention == "unidirectional":
            r, c = _grid(nb)
            same = same & (c <= r)
        layout[h] |= same.long()
        return layout

    def set_global_layout(self, h, layout):
        nb = layout.shape[1]
        ends = self.global_block_end_indices or [i + 1 for i in self.global_block_indices]
        for s, e in zip(self.global_block_indices, ends):
            if s < nb:
                e = min(e, nb)
                if self.horizontal_global_attention:
                    layout[h, s:e, :] = 1
                layout[h, (0 if self.attention == "bidirectional" else s):, s:e] = 1
        return layout

    def make_layout(self, seq_len):
        layout = self.setup_layout(seq_len)
        for h in range(self.num_layout_heads):
            layout = self.set_global_layout(h, self.set_local_layout(h, self.set_random_layout(h, layout)))
        return self.check_and_propagate_first_head_layout(layout)


class BigBirdSparsityConfig(SparsityConfig):

    def __init__(self, num_heads, block=16, different_layout_per_head=False, num_random_blocks=1,
                 num_sliding_window_blocks=3, num_global_blocks=1, attention="bidirectional"):
        super().__init__(num_heads, block, different_layout_per_head)
        if attention not in ("unidirectional", "bidirectional"):
            raise NotImplementedError('only "uni/bi-directional" attentions are supported for now!')
        self.num_random_blocks, self.num_sliding_window_blocks = num_random_blocks, num_sliding_window_blocks
        self.num_global_blocks, self.attention = num_global_blocks, attention

    def set_random_layout(self, h, layout):
        layout[h] |= self._random(layout.shape[1], self.num_random_blocks, self.attention == "unidirectional")
        return layout

    def set_sliding_window_layout(self, h, layout):
        layout[h] |= self._sliding_window(layout.shape[1], self.num_sliding_window_blocks)
        return layout

    def set_global_layout_itc(self, h, layout):
        nb = layout.shape[1]
        if nb < self.num_global_blocks:
            raise ValueError(f"Number of global blocks, {self.num_global_blocks}, must be smaller than overall number "
                             f"of blocks in a row, {nb}!")
        layout[h, :self.num_global_blocks, :] = 1
        layout[h, :, :self.num_global_blocks] = 1
        if self.attention == "unidirectional":
            layout[h] = torch.tril(layout[h])
        return layout

    def make_layout(self, seq_len):
        layout = self.setup_layout(seq_len)
        for h in range(self.num_layout_heads):
            layout = self.set_random_layout(h, layout)
            layout = self.set_sliding_window_layout(h, layout)
            layout = self.set_global_layout_itc(h, layout)
        return self.check_and_propagate_first_head_layout(layout)


class BSLongformerSparsityConfig(SparsityConfig):

    def __init__(self, num_heads, block=16, different_layout_per_head=False, num_sliding_window_blocks=3,
                 global_block_indices=(0, ), global_block_end_indices=None, attention="bidirectional"):
        super().__init__(num_heads, block, different_layout_per_head)
        if global_block_end_indices is not None:
            if len(global_block_indices) != len(global_block_end_indices):
                raise ValueError("Global block start and end index lists must have the same length!")
            for s, e in zip(global_block_indices, global_block_end_indices):
                if s >= e:
                    raise ValueError(f"Global block start index, {s}, must be smaller than global block end index, {e}!")
        self.num_sliding_window_blocks = num_sliding_window_blocks
        self.global_block_indices = list(global_block_indices)
        self.global_block_end_indices = list(global_block_end_indices) if global_block_end_indices is not None else None
        self.attention = attention

    def set_sliding_window_layout(self, h, layout):
        layout[h] |= self._sliding_window(layout.shape[1], self.num_sliding_window_blocks)
        return layout

    def set_global_layout(self, h, layout):
        nb = layout.shape[1]
        ends = self.global_block_end_indices or [i + 1 for i in self.global_block_indices]
        for s, e in zip(self.global_block_indices, ends):
            if s < nb:
                e = min(e, nb)
                layout[h, s:e, :] = 1
                layout[h, :, s:e] = 1
        if self.attention == "unidirectional":
            layout[h] = torch.tril(layout[h])
        return layout

    def make_layout(self, seq_len):
        layout = self.setup_layout(seq_len)
        for h in range(self.num_layout_heads):
            layout = self.set_global_layout(h, self.set_sliding_window_layout(h, layout))
        return self.check_and_propagate_first_head_layout(layout)


class LocalSlidingWindowSparsityConfig(SparsityConfig):

    def __init__(self, num_heads, block=16, num_sliding_window_blocks=3, attention="unidirectional"):
        super().__init__(num_heads, block)
        self.num_sliding_window_blocks, self.attention = num_sliding_window_blocks, attention

    def set_sliding_window_layout(self, h, layout):
        nb = layout.shape[1]
        win = self._sliding_window(nb, self.num_sliding_window_blocks)
        if self.attention == "unidirectional":
            win = torch.tril(win)
        layout[h] |= win
        return layout

    def make_layout(self, seq_len):
        layout = self.setup_layout(seq_len)
        for h in range(self.num_layout_heads):
            layout = self.set_sliding_window_layout(h, layout)
        return self.check_and_propagate_first_head_layout(layout)
