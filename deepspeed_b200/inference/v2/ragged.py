"""Ragged batching state: blocked KV-cache allocator, per-sequence descriptors, ragged batch wrapper.

Reference: ``inference/v2/ragged/`` (``BlockedAllocator`` blocked_allocator.py, ``DSSequenceDescriptor``
sequence_descriptor.py, ``DSStateManager`` manager.py, ``RaggedBatchWrapper`` ragged_wrapper.py, ``BlockedKVCache``
kv_cache.py)."""
from dataclasses import dataclass, field
from typing import Dict, List

import torch


class BlockedAllocator:
    """Free-list allocator of fixed-size KV blocks (linked list in a tensor, like the reference)."""

    def __init__(self, num_blocks):
        if num_blocks < 1:
            raise ValueError(f"Blocked KV-cache must have at least 1 block, provided {num_blocks}")
        self._num_blocks = num_blocks
        self._blocks = torch.arange(1, num_blocks + 1, dtype=torch.int32)
        self._head = 0
        self._free_blocks = num_blocks

    def allocate(self, num_blocks):
        if num_blocks > self._free_blocks:
            raise ValueError(f"Not enough free blocks in the KV-cache to allocate {num_blocks} blocks")
        out = torch.zeros(num_blocks, dtype=torch.int32)
        for i in range(num_blocks):
            out[i] = self._head
            self._head = int(self._blocks[self._head])
            self._blocks[int(out[i])] = -1
            self._free_blocks -= 1
        return out

    def free(self, blocks):
        if isinstance(blocks, int):
            blocks = [blocks]
        for b in (blocks.tolist() if torch.is_tensor(blocks) else blocks):
            if b < 0 or b >= self._num_blocks:
                raise ValueError(f"Invalid block {b} provided to free")
            if int(self._blocks[b]) != -1:
                raise ValueError(f"Block {b} is already free")
            self._blocks[b] = self._head
            self._head = b
            self._free_blocks += 1

    @property
    def free_blocks(self):
        return self._free_blocks


@dataclass
class DSSequenceDescriptor:
    uid: int
    seen_tokens: int = 0
    in_flight_tokens: int = 0
    blocks: List[int] = field(default_factory=list)

    @property
    def cur_allocated_blocks(self):
        return len(self.blocks)

    def pre_forward(self, n):
        self.in_flight_tokens = n

    def post_forward(self):
        self.seen_tokens += self.in_flight_tokens
        self.in_flight_tokens = 0


class BlockedKVCache:
    """[layers, num_blocks, 2 (K | V), block_size, kv_heads, head_dim]: the K (or V) plane of a block is one contiguous
    [block_size, kv_heads * head_dim] matrix (what the paged-attention kernels of ``csrc/ragged_ops.cu`` stream)."""

    def __init__(self, num_layers, num_blocks, block_size, kv_heads, head_dim, dtype, device):
        self.block_size = block_size
        self.cache = torch.zeros(num_layers, num_blocks, 2, block_size, kv_heads, head_dim, dtype=dtype, device=device)
        self.allocator = BlockedAllocator(num_blocks)

    def write(self, layer, seq: DSSequenceDescriptor, k, v, start):
        """k, v: [n, kv_heads, head_dim] appended at token positions start..start+n."""
        n = k.shape[0]
        pos = torch.arange(start, start + n, device=k.device)
        blk = torch.tensor(seq.blocks, device=k.device)[pos // self.block_size]
        off = pos % self.block_size
        self.cache[layer, blk, 0, off] = k
        self.cache[layer, blk, 1, off] = v

    def read(self, layer, seq: DSSequenceDescriptor, length):
        blk = torch.tensor(seq.blocks, device=self.cache.device)
        kv = self.cache[layer, blk]  # [n_blocks, 2, block_size, heads, dim]
        heads, dim = self.cache.shape[-2], self.cache.shape[-1]
        return kv[:, 0].reshape(-1, heads, dim)[:length], kv[:, 1].reshape(-1, heads, dim)[:length]


class DSStateManager:

    def __init__(self, kv_cache: BlockedKVCache, max_tracked_sequences=2048, max_ragged_batch_size=768,
                 max_ragged_sequence_count=512):
        self.kv = kv_cache
        self.seqs: Dict[int, DSSequenceDescriptor] = {}
        self.max_tracked_sequences = max_tracked_sequences
        self.max_ragged_batch_size, self.max_ragged_sequence_count = max_ragged_batch_size, max_ragged_sequence_count

    def get_or_create_sequence(self, uid):
        if uid not in self.seqs:
            if len(self.seqs) >= self.max_tracked_sequences:
                raise RuntimeError("too many tracked sequences")
            self.seqs[uid] = DSSequenceDescriptor(uid)
        return self.seqs[uid]

    def get_sequence(self, uid):
        return self.seqs.get(uid)

    def ensure_capacity(self, seq, new_tokens):
        need = -(-(seq.seen_tokens + new_tokens) // self.kv.block_size) - seq.cur_allocated_blocks
        if need > 0:
            seq.blocks += self.kv.allocator.allocate(need).tolist()

    def query(self, uid, new_tokens):
        """(tokens that can be scheduled, blocks needed) for admission control."""
        seq = self.seqs.get(uid) or DSSequenceDescriptor(uid)
        need = max(0, -(-(seq.seen_tokens + new_tokens) // self.kv.block_size) - seq.cur_allocated_blocks)
        return (new_tokens if need <= self.kv.allocator.free_blocks else 0), need

    def flush_sequence(self, uid):
        seq = self.seqs.pop(uid, None)
        if seq is not None and seq.blocks:
            self.kv.allocator.free(seq.blocks)

    @property
    def free_blocks(self):
        return self.kv.allocator.free_blocks

    @property
    def n_tracked_sequences(self):
        return len(self.seqs)


class RaggedBatchWrapper:
    """Flattened tokens of several sequences + per-sequence (start, length, seen) metadata."""

    def __init__(self):
        self.clear()

    def clear(self):
        self.tokens, self.seq_meta = [], []

    def insert_sequence(self, seq: DSSequenceDescriptor, tokens):
        start = sum(len(t) for t in self.tokens)
        self.tokens.append(tokens)
        self.seq_meta.append((seq, start, len(tokens), seq.seen_tokens))

    def finalize(self, device):
        return torch.cat(self.tokens).to(device)

    def device_metadata(self, device, max_blocks=None):
        """Per-token (sequence row, absolute position) and the block table of the batch, as int32 device tensors:
        the inputs of the ragged kernels (reference ``ragged_wrapper.py`` inflight descriptors + ``kv_ptrs``)."""
        tok_seq, tok_pos = [], []
        for row, (seq, start, n, seen) in enumerate(self.seq_meta):
            tok_seq += [row] * n
            tok_pos += list(range(seen, seen + n))
        width = max_blocks or max(1, max(len(seq.blocks) for seq, _, _, _ in self.seq_meta))
        table = torch.zeros(len(self.seq_meta), width, dtype=torch.int32)
        for row, (seq, _, _, _) in enumerate(self.seq_meta):
            if seq.blocks:
                table[row, :len(seq.blocks)] = torch.tensor(seq.blocks, dtype=torch.int32)
        host = torch.tensor([tok_seq, tok_pos], dtype=torch.int32)
        dev = host.to(device, non_blocking=True)
        return dev[0].contiguous(), dev[1].contiguous(), table.to(device, non_blocking=True)

    @property
    def current_tokens(self):
        return sum(len(t) for t in self.tokens)

    @property
    def current_sequences(self):
        return len(self.seq_meta)
