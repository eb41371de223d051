"""FastGen-style engine: ``put(uids, tokens)`` runs one ragged forward (prefill chunks and decode tokens mixed) over
a blocked KV cache and returns next-token logits per sequence.

Reference: ``inference/v2/engine_v2.py:30`` (``InferenceEngineV2.put/query/flush``), model implementations
``inference/v2/model_implementations/llama_v2`` (here: the framework's own Llama/Mixtral weights), ragged kernels
``inference/v2/kernels/ragged_ops`` (blocked rotary + KV copy, blocked flash). On a GPU the attention of the whole
ragged batch is two launches of ``csrc/ragged_ops.cu`` (RoPE + KV append, paged attention with split-KV); the
per-sequence SDPA over gathered blocks remains as the CPU / unsupported-shape path. Linear layers / norms are the
fused ops of this framework."""
import math
import os

import torch
import torch.nn.functional as F

from ...ops import fused
from ...ops import kernels as K
from .ragged import BlockedKVCache, DSStateManager, NativeRaggedState, RaggedBatchWrapper


class RaggedInferenceEngineConfig:
    """Flat keyword form, or the reference's nested form (``config_v2.py:30``): ``tensor_parallel`` / ``tp``
    ``{"tp_size": n}``, ``state_manager`` / ``manager`` ``{"max_tracked_sequences", "max_ragged_batch_size",
    "max_ragged_sequence_count", "max_context", "memory_config": {"mode": "reserve"|"allocate", "size": ...}}``,
    ``quantization`` ``{"quantization_mode": ...}``."""

    def __init__(self, kv_block_size=64, num_kv_blocks=1024, max_ragged_batch_size=768, max_ragged_sequence_count=512,
                 max_tracked_sequences=2048, tensor_parallel=None, tp=None, state_manager=None, manager=None,
                 quantization=None, dtype=None):
        from types import SimpleNamespace
        sm = dict(state_manager or manager or {})
        self.kv_block_size, self.num_kv_blocks = kv_block_size, num_kv_blocks
        self.max_ragged_batch_size = sm.get("max_ragged_batch_size", max_ragged_batch_size)
        self.max_ragged_sequence_count = sm.get("max_ragged_sequence_count", max_ragged_sequence_count)
        self.max_tracked_sequences = sm.get("max_tracked_sequences", max_tracked_sequences)
        self.max_context = sm.get("max_context", 8192)
        mem = dict(sm.get("memory_config", {}) or {})
        if str(mem.get("mode", "")).lower().endswith("allocate") and "size" in mem:
            self.num_kv_blocks = int(mem["size"])  # ALLOCATE mode: size is the number of KV blocks
        self.memory_config = SimpleNamespace(mode=mem.get("mode", "reserve"), size=mem.get("size", 1_000_000_000))
        tpd = tensor_parallel if tensor_parallel is not None else tp
        tp_size = tpd.get("tp_size", 1) if isinstance(tpd, dict) else getattr(tpd, "tp_size", 1) if tpd else 1
        self.tensor_parallel = SimpleNamespace(tp_size=int(tp_size))
        q = dict(quantization or {})
        self.quantization = SimpleNamespace(quantization_mode=q.get("quantization_mode"))
        self.state_manager = SimpleNamespace(max_tracked_sequences=self.max_tracked_sequences,
                                             max_ragged_batch_size=self.max_ragged_batch_size,
                                             max_ragged_sequence_count=self.max_ragged_sequence_count,
                                             max_context=self.max_context, memory_config=self.memory_config)
        self.dtype = dtype


class InferenceEngineV2:

    def __init__(self, model, config: RaggedInferenceEngineConfig = None):
        self.model = model.eval()
        self.cfg = model.cfg
        self.config = config or RaggedInferenceEngineConfig()
        p = next(model.parameters())
        self.device, self.dtype = p.device, p.dtype
        c = self.cfg
        kv = BlockedKVCache(c.num_hidden_layers, self.config.num_kv_blocks, self.config.kv_block_size,
                            c.num_key_value_heads, c.head_dim, self.dtype, self.device)
        self.state_manager = DSStateManager(kv, self.config.max_tracked_sequences, self.config.max_ragged_batch_size,
                                            self.config.max_ragged_sequence_count)
        self.batch = RaggedBatchWrapper()
        self._meta = None
        # C++ host runtime (allocator + sequence table + batch descriptors) whenever the native kernels serve this model
        self.native_state = None
        probe = torch.empty(0, dtype=self.dtype, device=self.device)
        if self._native(probe, c.num_attention_heads, c.num_key_value_heads, c.head_dim):
            try:
                self.native_state = NativeRaggedState(self.config.num_kv_blocks, self.config.kv_block_size,
                                                      self.config.max_tracked_sequences,
                                                      self.config.max_ragged_batch_size,
                                                      self.config.max_ragged_sequence_count)
            except Exception:  # noqa: BLE001 - host library unavailable: Python bookkeeping
                self.native_state = None
        self.generic = hasattr(model, "forward_tokens")  # model_implementations.GenericDecoder (any HF family)
        if not self.generic:
            self.cos, self.sin = fused.rope_tables(c.max_position_embeddings, c.head_dim, c.rope_theta, self.device,
                                                   c.rope_scaling)

    @property
    def _sm(self):
        return self.native_state if self.native_state is not None else self.state_manager

    @property
    def n_kv_cache_groups(self):
        """KV cache groups (one: every layer shares the block geometry)."""
        return 1

    def get_remaining_block_capacity(self, uid):
        """Tokens that still fit into the last KV block already allocated to sequence ``uid`` (0 if unknown)."""
        bs = self.config.kv_block_size
        if self.native_state is not None:
            seen = self.native_state.seen_tokens(uid)
            if seen < 0:
                return 0
        else:
            seq = self.state_manager.get_sequence(uid)
            if seq is None:
                return 0
            seen = seq.seen_tokens
        return (-seen) % bs

    def serialize(self, save_path):
        """Write the model so that ``build_engine_from_ds_checkpoint`` can rebuild the engine without the original
        HuggingFace files (reference ``engine_v2.py:251``): parameters, metadata naming the policy, pickled config."""
        import json
        import os
        import pickle
        from .engine_factory import DS_MODEL_CONFIG, make_metadata_filename, make_param_filename
        hf_config = getattr(self.model, "hf_config", None)
        if hf_config is None:
            raise RuntimeError("serialize() needs a model built from a HuggingFace config (build_hf_engine)")
        os.makedirs(save_path, exist_ok=True)
        tp = self.config.tensor_parallel.tp_size
        torch.save({k: v.detach().cpu() for k, v in self.model.state_dict().items()},
                   make_param_filename(save_path, 0, tp))
        meta = {"policy": hf_config.model_type,
                "params": {k: list(v.shape) for k, v in self.model.state_dict().items()}}
        json.dump(meta, open(make_metadata_filename(save_path, 0, tp), "w"))
        with open(os.path.join(save_path, DS_MODEL_CONFIG), "wb") as f:
            pickle.dump(hf_config, f)

    @property
    def free_blocks(self):
        return self._sm.free_blocks

    def query(self, uid, max_request_tokens, max_request_blocks=None):
        return self._sm.query(uid, max_request_tokens)

    def can_schedule(self, uids, lengths):
        total_blocks = 0
        if sum(lengths) > self.config.max_ragged_batch_size or len(uids) > self.config.max_ragged_sequence_count:
            return False
        for uid, n in zip(uids, lengths):
            ok, need = self._sm.query(uid, n)
            if ok == 0:
                return False
            total_blocks += need
        return total_blocks <= self.free_blocks

    def flush(self, uid):
        self._sm.flush_sequence(uid)

    @torch.no_grad()
    def put(self, batch_uids, batch_tokens, do_checks=True):
        """Returns logits [n_sequences, vocab] for the last token of each sequence chunk."""
        if do_checks and not self.can_schedule(batch_uids, [len(t) for t in batch_tokens]):
            raise RuntimeError("ragged batch cannot be scheduled (KV blocks / batch limits)")
        self.batch.clear()
        if self.native_state is not None:
            # one C++ call builds every descriptor of this forward; one pinned copy ships them
            toks = [torch.as_tensor(t, dtype=torch.long) for t in batch_tokens]
            lens = [int(t.numel()) for t in toks]
            pmin = self.PREFILL_MIN_TOKENS if self.config.kv_block_size == 128 else (1 << 30)
            dmax = 1 if self.config.kv_block_size == 128 else 0  # decode class only with 128-token blocks
            meta = self.native_state.schedule(batch_uids, lens, pmin, self.device)
            if dmax == 0 and meta["decode"] is not None:  # fold single-token sequences into the SIMT group
                meta["decode"] = None
                meta["rest"] = None
            for grp in ("prefill", "decode"):
                if meta[grp] is not None:
                    meta[grp]["table"] = meta["table"].index_select(0, meta[grp]["seq_rows"].long()).contiguous()
            if meta["decode"] is not None:
                meta["decode"]["rows"] = meta["decode"]["rows"].long()
            if meta["prefill"] is None and meta["decode"] is None:
                meta["rest"] = None  # everything goes through the SIMT kernel with the whole-batch arrays
            self._meta = meta
            ids = torch.cat(toks).to(self.device, non_blocking=True)
            ends, acc = [], 0
            for n in lens:
                acc += n
                ends.append(acc - 1)
            last = torch.tensor(ends, device=self.device)
            return self._forward(ids, meta["tok_pos"].long(), last)
        for uid, toks in zip(batch_uids, batch_tokens):
            seq = self.state_manager.get_or_create_sequence(uid)
            toks = torch.as_tensor(toks, dtype=torch.long)
            self.state_manager.ensure_capacity(seq, len(toks))
            seq.pre_forward(len(toks))
            self.batch.insert_sequence(seq, toks)
        ids = self.batch.finalize(self.device)
        self._meta = None
        pos = torch.cat([torch.arange(seen, seen + n, device=self.device) for (_, _, n, seen) in self.batch.seq_meta])
        last = torch.tensor([s + n - 1 for (_, s, n, _) in self.batch.seq_meta], device=self.device)
        logits = self._forward(ids, pos, last)
        for seq, _, _, _ in self.batch.seq_meta:
            seq.post_forward()
        return logits

    # ---- native ragged kernels --------------------------------------------------------------------
    def _native(self, q, nh, nkv, hd):
        return (os.environ.get("DSB200_RAGGED_NATIVE", "1") == "1" and q.is_cuda
                and K.ragged_attention_supported(q, hd, nh, nkv))

    # With head_dim 128 and 128-token KV blocks the attention of a ragged batch runs on the tcgen05 kernel: single-token
    # (decode) sequences through its split-KV decode mode, chunks of at least PREFILL_MIN_TOKENS tokens through its
    # paged prefill mode; what is left (short chunks, other head dims / block sizes) goes to the SIMT paged kernel.
    PREFILL_MIN_TOKENS = 128

    def _batch_meta(self):
        if self._meta is None:
            dev = self.device
            sm = self.batch.seq_meta
            tok_seq, tok_pos, table = self.batch.device_metadata(dev)
            meta = dict(tok_seq=tok_seq, tok_pos=tok_pos, table=table, prefill=None, decode=None, rest=None,
                        max_ctx=max(seen + n for (_, _, n, seen) in sm))
            if self.config.kv_block_size == 128:
                long_rows = [r for r, (_, _, n, _) in enumerate(sm) if n >= self.PREFILL_MIN_TOKENS]
                dec_rows = [r for r, (_, _, n, _) in enumerate(sm) if n == 1]
                rest = [r for r, (_, _, n, _) in enumerate(sm) if 1 < n < self.PREFILL_MIN_TOKENS]
                if long_rows:
                    ints = torch.tensor([[sm[r][1] for r in long_rows], [sm[r][2] for r in long_rows],
                                         [sm[r][3] + sm[r][2] for r in long_rows]], dtype=torch.int32).to(dev)
                    meta["prefill"] = dict(cu_q=ints[0].contiguous(), q_len=ints[1].contiguous(),
                                           kv_len=ints[2].contiguous(),
                                           table=table[torch.tensor(long_rows, device=dev)].contiguous(),
                                           max_q=max(sm[r][2] for r in long_rows))
                if dec_rows:
                    ints = torch.tensor([[sm[r][1] for r in dec_rows], [sm[r][3] + 1 for r in dec_rows]],
                                        dtype=torch.int32).to(dev)
                    meta["decode"] = dict(rows=ints[0].long(), kv_len=ints[1].contiguous(),
                                          table=table[torch.tensor(dec_rows, device=dev)].contiguous(),
                                          max_ctx=max(sm[r][3] + 1 for r in dec_rows), all=len(dec_rows) == len(sm))
                if rest and (long_rows or dec_rows):
                    rows, rseq, rpos = [], [], []
                    for r in rest:
                        _, start, n, seen = sm[r]
                        rows += list(range(start, start + n))
                        rseq += [r] * n
                        rpos += list(range(seen, seen + n))
                    t = torch.tensor([rows, rseq, rpos], dtype=torch.int32).to(dev)
                    meta["rest"] = dict(rows=t[0].contiguous(), seq=t[1].contiguous(), pos=t[2].contiguous(),
                                        max_ctx=max(sm[r][3] + sm[r][2] for r in rest))
            self._meta = meta
        return self._meta

    def _attend_native(self, li, q, k, v, cos=None, sin=None, rotary_dim=0):
        """q [T, Hq, D] (rotated in place when ``rotary_dim``), k / v [T, Hkv, D] -> [T, Hq * D]."""
        from ...ops import attention as A
        m = self._batch_meta()
        kv_layer = self.state_manager.kv.cache[li]
        scale = 1.0 / math.sqrt(q.shape[-1])
        K.ragged_kv_append_rope(q, k, v, kv_layer, cos, sin, m["tok_seq"], m["tok_pos"], m["table"], rotary_dim)
        pf, dc, rest = m["prefill"], m["decode"], m["rest"]
        if (pf is None and dc is None) or not A.paged_prefill_supported(q, kv_layer):
            o = K.ragged_paged_attention(q, kv_layer, m["tok_seq"], m["tok_pos"], m["table"], scale, m["max_ctx"])
            return o.reshape(q.shape[0], -1)
        o = torch.empty((q.shape[0], q.shape[1], q.shape[2]), dtype=q.dtype, device=q.device)
        if pf is not None:
            A.flash_attention_paged_prefill(q, kv_layer, pf["cu_q"], pf["q_len"], pf["kv_len"], pf["table"], pf["max_q"],
                                            scale, out=o)
        if dc is not None:
            qd = q.contiguous() if dc["all"] else q.index_select(0, dc["rows"])
            od = A.flash_decode_paged(qd, kv_layer, dc["kv_len"], dc["table"], dc["max_ctx"], scale,
                                      out=o if dc["all"] else None)
            if not dc["all"]:
                o.index_copy_(0, dc["rows"], od)
        if rest is not None:
            K.ragged_paged_attention(q, kv_layer, rest["seq"], rest["pos"], m["table"], scale, rest["max_ctx"], out=o,
                                     tok_row=rest["rows"])
        return o.reshape(q.shape[0], -1)

    def _attend(self, li, q, k, v):
        """Blocked-KV attention for the current ragged batch: q/k/v are [T, heads, head_dim] with positions applied."""
        nh, nkv = q.shape[1], k.shape[1]
        if self._native(q, nh, nkv, q.shape[2]):
            return self._attend_native(li, q, k, v)
        outs = []
        for (seq, start, n, seen) in self.batch.seq_meta:
            self.state_manager.kv.write(li, seq, k[start:start + n], v[start:start + n], seen)
            kk, vv = self.state_manager.kv.read(li, seq, seen + n)
            qq = q[start:start + n].transpose(0, 1).unsqueeze(0)
            kk = kk.transpose(0, 1).repeat_interleave(nh // nkv, 0).unsqueeze(0)
            vv = vv.transpose(0, 1).repeat_interleave(nh // nkv, 0).unsqueeze(0)
            mask = (torch.arange(seen + n, device=self.device)[None, :] <=
                    (torch.arange(n, device=self.device) + seen)[:, None])
            o = F.scaled_dot_product_attention(qq, kk, vv, attn_mask=mask)
            outs.append(o[0].transpose(0, 1).reshape(n, -1))
        return torch.cat(outs, 0)

    def _forward(self, ids, pos, last):
        m = self.model
        if self.generic:
            h = m.forward_tokens(ids, pos, self._attend)
            return m.lm_head(h.index_select(0, last))
        inner = getattr(m, "model", m)
        layers = inner.layers
        h = inner.embed_tokens(ids)  # [T, H]
        nh, nkv, hd = self.cfg.num_attention_heads, self.cfg.num_key_value_heads, self.cfg.head_dim
        for li, layer in enumerate(layers):
            x = layer.input_layernorm(h)
            attn = layer.self_attn
            qkv = attn.qkv_proj(x)
            lnh, lnkv = attn.nh, attn.nkv
            q, k, v = qkv.split([lnh * hd, lnkv * hd, lnkv * hd], dim=-1)
            if self._native(qkv, lnh, lnkv, hd):
                T = qkv.shape[0]
                a = attn.o_proj(self._attend_native(li, q.view(T, lnh, hd), k.view(T, lnkv, hd), v.view(T, lnkv, hd),
                                                    self.cos, self.sin, hd))
                x2, h = layer.post_attention_layernorm(a, residual=h)
                ffn = layer.mlp if hasattr(layer, "mlp") else layer.block_sparse_moe
                h = h + ffn(x2.unsqueeze(0))[0] if not hasattr(layer, "mlp") else h + ffn(x2)
                continue
            q = fused.apply_rope(q.reshape(1, -1, lnh, hd), self.cos, self.sin, pos.view(1, -1))[0]
            k = fused.apply_rope(k.reshape(1, -1, lnkv, hd), self.cos, self.sin, pos.view(1, -1))[0]
            v = v.reshape(-1, lnkv, hd)
            outs = []
            for (seq, start, n, seen) in self.batch.seq_meta:
                self.state_manager.kv.write(li, seq, k[start:start + n], v[start:start + n], seen)
                kk, vv = self.state_manager.kv.read(li, seq, seen + n)
                qq = q[start:start + n].transpose(0, 1).unsqueeze(0)  # [1, h, n, d]
                rep = lnh // lnkv
                kk = kk.transpose(0, 1).repeat_interleave(rep, 0).unsqueeze(0)
                vv = vv.transpose(0, 1).repeat_interleave(rep, 0).unsqueeze(0)
                mask = (torch.arange(seen + n, device=self.device)[None, :] <=
                        (torch.arange(n, device=self.device) + seen)[:, None])
                o = F.scaled_dot_product_attention(qq, kk, vv, attn_mask=mask)
                outs.append(o[0].transpose(0, 1).reshape(n, lnh * hd))
            a = attn.o_proj(torch.cat(outs, 0))
            x2, h = layer.post_attention_layernorm(a, residual=h)
            ffn = layer.mlp if hasattr(layer, "mlp") else layer.block_sparse_moe
            h = h + ffn(x2.unsqueeze(0))[0] if not hasattr(layer, "mlp") else h + ffn(x2)
        h = inner.norm(h)
        head = m.lm_head
        return head(h.index_select(0, last))


@torch.no_grad()
def greedy_generate(model, input_ids, max_new_tokens=16, **kw):
    """Greedy decoding through the ragged engine (one sequence per row of ``input_ids``)."""
    eng = InferenceEngineV2(model)
    out = [list(map(int, row)) for row in input_ids.tolist()]
    uids = list(range(len(out)))
    logits = eng.put(uids, [torch.tensor(r) for r in out])
    for _ in range(max_new_tokens):
        nxt = logits.argmax(-1).tolist()
        for r, t in zip(out, nxt):
            r.append(int(t))
        logits = eng.put(uids, [torch.tensor([t]) for t in nxt])
    for u in uids:
        eng.flush(u)
    return torch.tensor(out)
