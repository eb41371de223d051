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


class NativeRaggedState:
    """Blocked-KV allocator + sequence table + batch descriptor builder in C++ (``csrc/cpu/ragged_host.cpp``): one call
    per forward produces every integer array the ragged kernels consume, in one pinned buffer -> one host-to-device copy.
    Reference: ``ragged/manager.py`` + ``ragged_wrapper.py`` + ``ragged/csrc/fast_host_buffer.cu`` +
    ``kernels/ragged_ops/atom_builder``."""

    def __init__(self, num_blocks, block_size, max_tracked_sequences=2048, max_tokens=768, max_sequences=512,
                 max_blocks_per_seq=None):
        import ctypes
        from ...ops import cpu_lib
        self.L = cpu_lib.lib()
        if not hasattr(self.L, "dsb200_ragged_create"):
            raise RuntimeError("libdsb200_cpu.so was built without the ragged host runtime")
        self.h = ctypes.c_void_p(self.L.dsb200_ragged_create(num_blocks, block_size, max_tracked_sequences))
        self.block_size, self.num_blocks = block_size, num_blocks
        width = max_blocks_per_seq or num_blocks
        self.cap = 5 * max_tokens + max_sequences * (width + 8) + 64
        self.buf = torch.empty(self.cap, dtype=torch.int32)
        if torch.cuda.is_available():
            self.buf = self.buf.pin_memory()
        self.header = torch.zeros(16, dtype=torch.int32)
        self._ct = ctypes

    def __del__(self):
        try:
            self.L.dsb200_ragged_destroy(self.h)
        except Exception:  # noqa: BLE001
            pass

    @property
    def free_blocks(self):
        return int(self.L.dsb200_ragged_free_blocks(self.h))

    @property
    def n_tracked_sequences(self):
        return int(self.L.dsb200_ragged_tracked(self.h))

    def seen_tokens(self, uid):
        return int(self.L.dsb200_ragged_seen(self.h, int(uid)))

    def query(self, uid, new_tokens):
        need = int(self.L.dsb200_ragged_blocks_needed(self.h, int(uid), int(new_tokens)))
        return (new_tokens if need <= self.free_blocks else 0), need

    def blocks(self, uid):
        out = torch.empty(self.num_blocks, dtype=torch.int32)
        n = int(self.L.dsb200_ragged_blocks(self.h, int(uid), self._ct.c_void_p(out.data_ptr()), out.numel()))
        return out[:max(n, 0)].tolist()

    def flush_sequence(self, uid):
        rc = int(self.L.dsb200_ragged_flush(self.h, int(uid)))
        if rc != 0:
            raise ValueError(f"KV block bookkeeping corrupted while flushing sequence {uid} (code {rc})")

    def schedule(self, uids, lengths, prefill_min_tokens, device, commit=True):
        """-> dict of int32 device tensors + host scalars describing this forward (see ragged_host.cpp)."""
        n = len(uids)
        u = torch.tensor([int(x) for x in uids], dtype=torch.int64)
        ln = torch.tensor([int(x) for x in lengths], dtype=torch.int32)
        P = self._ct.c_void_p
        rc = int(self.L.dsb200_ragged_schedule(self.h, n, P(u.data_ptr()), P(ln.data_ptr()), int(prefill_min_tokens),
                                               P(self.buf.data_ptr()), self.cap, P(self.header.data_ptr()),
                                               1 if commit else 0))
        if rc == -1:
            raise ValueError("Not enough free blocks in the KV-cache for this ragged batch")
        if rc == -2:
            raise RuntimeError("too many tracked sequences")
        if rc == -3:
            raise RuntimeError("ragged batch larger than the descriptor buffer (raise max_ragged_batch_size)")
        hd = self.header.tolist()
        T, n_seq, width, n_p, max_q, n_d, ctx_d, T_r, ctx_r, max_ctx, o_p, o_d, o_r, total = hd[:14]
        dev = self.buf[:total].to(device, non_blocking=True)
        meta = dict(T=T, tok_seq=dev[:T], tok_pos=dev[T:2 * T], table=dev[2 * T:2 * T + n_seq * width].view(n_seq, width),
                    max_ctx=max_ctx, prefill=None, decode=None, rest=None)
        if n_p:
            pf = dev[o_p:o_p + 4 * n_p].view(4, n_p)
            meta["prefill"] = dict(cu_q=pf[0], q_len=pf[1], kv_len=pf[2], seq_rows=pf[3], max_q=max_q)
        if n_d:
            dc = dev[o_d:o_d + 3 * n_d].view(3, n_d)
            meta["decode"] = dict(rows=dc[0], kv_len=dc[1], seq_rows=dc[2], max_ctx=ctx_d, all=(n_d == n_seq and T == n_d))
        if T_r:
            rs = dev[o_r:o_r + 3 * T_r].view(3, T_r)
            meta["rest"] = dict(rows=rs[0], seq=rs[1], pos=rs[2], max_ctx=ctx_r)
        return meta
