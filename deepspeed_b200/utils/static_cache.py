"""Static KV cache for CUDA-graph decode of HuggingFace models (reference ``utils/static_cache.py``).

A captured graph freezes every host-side counter, so the position new keys/values are written at must come from a
device tensor at a fixed address: ``set_write_position(pos)`` stores a scalar ``torch.long`` tensor BY REFERENCE, the
caller advances it in place (``pos.add_(1)``) before each replay and every layer writes at ``pos + arange(new)``.

All layers live in one allocation ``[layers, 2, batch, kv_heads, max_len, head_dim]`` (one address range to keep
static, one memset to reset); ``layers[i].keys`` / ``.values`` are views of it. Duck-types ``transformers.StaticCache``.
"""
import torch


class DeepSpeedStaticLayer:
    is_compileable = True
    is_sliding = False

    def __init__(self, max_cache_len):
        self.max_cache_len = max_cache_len
        self.keys = None
        self.values = None
        self.is_initialized = False
        self._write_position = None
        self._arange = None

    def set_write_position(self, pos):
        self._write_position = pos

    def _bind(self, keys, values):
        self.keys, self.values = keys, values
        self.dtype, self.device = keys.dtype, keys.device
        self.max_batch_size, self.num_heads = keys.shape[:2]
        self.k_head_dim, self.v_head_dim = keys.shape[-1], values.shape[-1]
        self._arange = torch.arange(self.max_cache_len, device=self.device)
        self.is_initialized = True

    def lazy_initialization(self, key_states, value_states):
        b, h = key_states.shape[:2]
        k = torch.zeros((b, h, self.max_cache_len, key_states.shape[-1]), dtype=key_states.dtype,
                        device=key_states.device)
        v = torch.zeros((b, h, self.max_cache_len, value_states.shape[-1]), dtype=value_states.dtype,
                        device=value_states.device)
        self._bind(k, v)

    def update(self, key_states, value_states, *args, **kwargs):
        if not self.is_initialized:
            self.lazy_initialization(key_states, value_states)
        n = key_states.shape[-2]
        where = self._arange[:n]
        if self._write_position is not None:
            where = where + self._write_position
        self.keys.index_copy_(2, where, key_states)
        self.values.index_copy_(2, where, value_states)
        return self.keys, self.values

    def get_mask_sizes(self, query_length):
        return self.max_cache_len, 0

    def get_seq_length(self):
        if not self.is_initialized or self._write_position is None:
            return 0
        return self._write_position + 1

    def get_max_cache_shape(self):
        return self.max_cache_len

    def reset(self):
        if self.is_initialized:
            self.keys.zero_()
            self.values.zero_()

    def reorder_cache(self, beam_idx):
        if self.is_initialized:
            idx = beam_idx.to(self.keys.device)
            self.keys.copy_(self.keys.index_select(0, idx))  # in place: the addresses must stay what the graph saw
            self.values.copy_(self.values.index_select(0, idx))


class DeepSpeedStaticCache:

    def __init__(self, config, batch_size=1, max_cache_len=4096, device=None, dtype=None):
        self.config = config
        text = getattr(config, "text_config", config)
        n_layers = getattr(text, "num_hidden_layers", 1)
        self._layers = [DeepSpeedStaticLayer(max_cache_len) for _ in range(n_layers)]
        self._max_cache_len = max_cache_len
        self._write_position = None
        self._storage = None
        if dtype is not None and device is not None and batch_size > 0:
            heads = getattr(text, "num_attention_heads", 1)
            kv_heads = getattr(text, "num_key_value_heads", None) or heads
            head_dim = getattr(text, "head_dim", None) or getattr(text, "hidden_size", heads) // heads
            self.early_initialization(batch_size, kv_heads, head_dim, dtype, device)

    @property
    def layers(self):
        return self._layers

    def set_write_position(self, pos):
        self._write_position = pos
        for layer in self._layers:
            layer.set_write_position(pos)

    def early_initialization(self, batch_size, num_heads, head_dim, dtype, device):
        self._storage = torch.zeros((len(self._layers), 2, batch_size, num_heads, self._max_cache_len, head_dim),
                                    dtype=dtype, device=device)
        for i, layer in enumerate(self._layers):
            layer._bind(self._storage[i, 0], self._storage[i, 1])

    def update(self, key_states, value_states, layer_idx, *args, **kwargs):
        if layer_idx >= len(self._layers):
            raise IndexError(f"layer_idx {layer_idx} out of range (cache has {len(self._layers)} layers)")
        return self._layers[layer_idx].update(key_states, value_states, *args, **kwargs)

    def get_seq_length(self, layer_idx=0):
        return 0 if layer_idx >= len(self._layers) else self._layers[layer_idx].get_seq_length()

    def get_max_cache_shape(self, layer_idx=0):
        return self._max_cache_len

    def get_mask_sizes(self, query_length, layer_idx=0):
        return self._max_cache_len, 0

    def reset(self):
        if self._storage is not None:
            self._storage.zero_()
        else:
            for layer in self._layers:
                layer.reset()

    def __len__(self):
        return len(self._layers)
