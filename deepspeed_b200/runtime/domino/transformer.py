"""Domino: hide tensor-parallel communication by splitting every batch into two micro-batches inside the layer, so
the all-reduce of one overlaps the compute of the other (reference ``runtime/domino/transformer.py:250-605``).

Forward, per layer:   attn(mb0) -> [async AR mb0] -> attn(mb1) -> [async AR mb1] -> wait mb0 -> mlp(mb0) -> [async AR]
                      -> wait mb1 -> mlp(mb1) -> [async AR] -> wait both
Backward overlap comes from :class:`DominoAsyncColumnParallelLinearImpl` (async input-gradient all-reduce) and
:class:`NoOper`, an identity whose backward waits on the handle parked by the layer that produced the gradient."""
import enum
import math

import torch
import torch.nn.functional as F

from ... import comm as dist
from .async_linear import DominoAsyncColumnParallelLinear, RowParallelLinearNoComm


class LayerType(enum.Enum):
    encoder = 1
    decoder = 2


class AttnType(enum.Enum):
    self_attn = 1
    cross_attn = 2


class AttnMaskType(enum.Enum):
    padding = 1
    causal = 2


class ModelType(enum.Enum):
    encoder_or_decoder = 1
    encoder_and_decoder = 2


class DominoUtil:
    BATCH_0 = "BATCH0"
    BATCH_1 = "BATCH1"
    HANDLE_DIC = {"BATCH0": None, "BATCH1": None}


class DominoModule(torch.nn.Module):
    """Base class for all modules of a Domino transformer."""


def _Wait_bwd_comm(input_, dic_, h_id):
    return NoOper.apply(input_, dic_, h_id)


class NoOper(torch.autograd.Function):

    @staticmethod
    def forward(ctx, input_, handle_dic, h_id):
        ctx.handle_dic, ctx.h_id = handle_dic, h_id
        return input_

    @staticmethod
    def backward(ctx, grad_output):
        handle = ctx.handle_dic.get(ctx.h_id)
        if handle is not None:
            handle.wait()
            ctx.handle_dic[ctx.h_id] = None
        return grad_output, None, None


class CoreAttention(DominoModule):

    def __init__(self, config, tp_world_size, attn_mask_type=AttnMaskType.causal):
        super().__init__()
        self.attn_mask_type = attn_mask_type
        self.hidden_size_per_partition = config.hidden_size // tp_world_size
        self.attention_dropout_rate = getattr(config, "attention_dropout", 0.0)

    def forward(self, query_layer, key_layer, value_layer, attention_mask):
        # [b, heads, s, d]
        p = self.attention_dropout_rate if self.training else 0.0
        ctx = F.scaled_dot_product_attention(query_layer, key_layer, value_layer, attn_mask=None, dropout_p=p,
                                             is_causal=(self.attn_mask_type == AttnMaskType.causal))
        b, h, s, d = ctx.shape
        return ctx.permute(0, 2, 1, 3).reshape(b, s, h * d)


class ShardedAttention(DominoModule):
    """TP-sharded self attention: async column-parallel QKV, row-parallel dense without communication."""

    def __init__(self, config, tp_group, tp_world_size, attn_mask_type=AttnMaskType.causal, dtype=torch.float32,
                 device=None):
        super().__init__()
        self.tp_world_size = tp_world_size
        self.heads_per_partition = config.num_attention_heads // tp_world_size
        self.head_dim = config.hidden_size // config.num_attention_heads
        self.query_key_value = DominoAsyncColumnParallelLinear(config.hidden_size, 3 * config.hidden_size, tp_group,
                                                               tp_world_size=tp_world_size, dtype=dtype, device=device)
        self.core_attention = CoreAttention(config, tp_world_size, attn_mask_type)
        self.dense = RowParallelLinearNoComm(config.hidden_size, config.hidden_size, skip_bias_add=True,
                                             tp_world_size=tp_world_size, dtype=dtype, device=device)

    def forward(self, hidden_states, attention_mask, micro_batch_num, rotary_pos_emb=None):
        mixed, _ = self.query_key_value(hidden_states, DominoUtil.HANDLE_DIC, micro_batch_num)
        b, s, _ = mixed.shape
        qkv = mixed.view(b, s, self.heads_per_partition, 3 * self.head_dim).permute(0, 2, 1, 3)
        q, k, v = qkv.split(self.head_dim, dim=-1)
        if rotary_pos_emb is not None:
            cos, sin = rotary_pos_emb
            q, k = _rope(q, cos, sin), _rope(k, cos, sin)
        ctx = self.core_attention(q, k, v, attention_mask)
        return self.dense(ctx)


def _rope(x, cos, sin):
    x1, x2 = x.chunk(2, dim=-1)
    return torch.cat([x1 * cos - x2 * sin, x2 * cos + x1 * sin], dim=-1)


class bias_dropout_add(torch.nn.Module):

    def __init__(self, prob: float):
        super().__init__()
        self.dropout = torch.nn.Dropout(prob)

    def forward(self, x, bias, residual):
        if bias is not None:
            x = x + bias
        return residual + self.dropout(x)


class DominoTransformerLayer(DominoModule):

    def __init__(self, config, mpu=None, apply_rotary_pos_emb=None, layer_number=1, layer_type=LayerType.encoder,
                 self_attn_mask_type=AttnMaskType.causal, drop_path_rate=0., tp_group=None, dtype=torch.float32,
                 device=None):
        super().__init__()
        self.layer_number = layer_number
        if tp_group is None and mpu is not None:
            tp_group = mpu.get_tensor_model_parallel_group()
        self.tp_group = tp_group
        self.tp_world = dist.get_world_size(tp_group) if (dist.is_initialized() and tp_group is not None) else \
            (mpu.get_tensor_model_parallel_world_size() if mpu is not None else 1)
        h = config.hidden_size
        ffn = getattr(config, "ffn_hidden_size", 4 * h)
        eps = getattr(config, "layernorm_epsilon", 1e-5)
        self.input_layernorm = torch.nn.LayerNorm(h, eps=eps, dtype=dtype, device=device)
        self.self_attention = ShardedAttention(config, tp_group, self.tp_world, self_attn_mask_type, dtype, device)
        self.attn_bda = bias_dropout_add(getattr(config, "hidden_dropout", 0.0))
        self.post_attention_layernorm = torch.nn.LayerNorm(h, eps=eps, dtype=dtype, device=device)
        self.dense_h_to_4h = DominoAsyncColumnParallelLinear(h, ffn, tp_group, tp_world_size=self.tp_world,
                                                             dtype=dtype, device=device)
        self.dense_4h_to_h = RowParallelLinearNoComm(ffn, h, skip_bias_add=True, tp_world_size=self.tp_world,
                                                     dtype=dtype, device=device)
        self.mlp_bda = bias_dropout_add(getattr(config, "hidden_dropout", 0.0))

    def _ar(self, t):
        if self.tp_world > 1:
            return dist.all_reduce(t, group=self.tp_group, async_op=True)
        return None

    def forward(self, hidden_states, attention_mask=None, rotary_pos_emb=None):
        """``hidden_states``: [b, s, h] with an even batch, processed as two micro-batches."""
        x0, x1 = hidden_states.chunk(2, dim=0)
        W0 = lambda t: _Wait_bwd_comm(t, DominoUtil.HANDLE_DIC, DominoUtil.BATCH_0)  # noqa: E731
        W1 = lambda t: _Wait_bwd_comm(t, DominoUtil.HANDLE_DIC, DominoUtil.BATCH_1)  # noqa: E731
        m0 = attention_mask.chunk(2, dim=0)[0] if attention_mask is not None and attention_mask.shape[0] > 1 else attention_mask
        m1 = attention_mask.chunk(2, dim=0)[1] if attention_mask is not None and attention_mask.shape[0] > 1 else attention_mask
        # ---- attention: the all-reduce of micro-batch 0 overlaps the attention compute of micro-batch 1
        a0, b0 = self.self_attention(W0(self.input_layernorm(x0)), m0, DominoUtil.BATCH_0, rotary_pos_emb)
        h0 = self._ar(a0)
        a1, b1 = self.self_attention(W1(self.input_layernorm(x1)), m1, DominoUtil.BATCH_1, rotary_pos_emb)
        h1 = self._ar(a1)
        if h0 is not None:
            h0.wait()
        r0 = self.attn_bda(a0, b0, x0)
        # ---- MLP of micro-batch 0 overlaps the attention all-reduce of micro-batch 1
        u0, _ = self.dense_h_to_4h(W0(self.post_attention_layernorm(r0)), DominoUtil.HANDLE_DIC, DominoUtil.BATCH_0)
        o0, ob0 = self.dense_4h_to_h(F.gelu(u0))
        g0 = self._ar(o0)
        if h1 is not None:
            h1.wait()
        r1 = self.attn_bda(a1, b1, x1)
        u1, _ = self.dense_h_to_4h(W1(self.post_attention_layernorm(r1)), DominoUtil.HANDLE_DIC, DominoUtil.BATCH_1)
        o1, ob1 = self.dense_4h_to_h(F.gelu(u1))
        g1 = self._ar(o1)
        if g0 is not None:
            g0.wait()
        y0 = self.mlp_bda(o0, ob0, r0)
        if g1 is not None:
            g1.wait()
        y1 = self.mlp_bda(o1, ob1, r1)
        return torch.cat([y0, y1], dim=0)


class DominoTransformer(DominoModule):

    def __init__(self, config, model_type=ModelType.encoder_or_decoder, mpu=None, apply_rotary_pos_emb=None,
                 layer_type=LayerType.encoder, self_attn_mask_type=AttnMaskType.causal, post_layer_norm=True,
                 pre_process=True, post_process=True, drop_path_rate=0.0, tp_group=None, dtype=torch.float32,
                 device=None):
        super().__init__()
        self.post_layer_norm, self.post_process = post_layer_norm, post_process
        self.layers = torch.nn.ModuleList([
            DominoTransformerLayer(config, mpu, apply_rotary_pos_emb, i + 1, layer_type, self_attn_mask_type,
                                   tp_group=tp_group, dtype=dtype, device=device)
            for i in range(config.num_layers)
        ])
        if post_process and post_layer_norm:
            self.final_layernorm = torch.nn.LayerNorm(config.hidden_size, eps=getattr(config, "layernorm_epsilon",
                                                                                       1e-5), dtype=dtype,
                                                      device=device)

    def forward(self, hidden_states, attention_mask=None, rotary_pos_emb=None):
        for layer in self.layers:
            hidden_states = layer(hidden_states, attention_mask, rotary_pos_emb)
        if self.post_process and self.post_layer_norm:
            hidden_states = self.final_layernorm(hidden_states)
        return hidden_states
