"""v1 MoE inference layer (reference ``ops/transformer/inference/moe_inference.py``): the attention block of
:class:`DeepSpeedTransformerInference` followed by a top-k gated mixture of expert MLPs (optionally the residual-MoE form
that adds a dense MLP weighted by a learned 2-way coefficient)."""
import torch
import torch.nn as nn
import torch.nn.functional as F

from .... import comm as dist
from .config import DeepSpeedInferenceConfig
from .ds_transformer import DeepSpeedSelfAttention, _empty, _norm


class DeepSpeedMoEInferenceConfig(DeepSpeedInferenceConfig):

    def __init__(self, hidden_size=-1, intermediate_size=-1, heads=-1, num_hidden_layers=-1, layer_norm_eps=1e-12,
                 local_rank=-1, mp_size=1, fp16=False, bf16=False, q_int8=False, pre_layer_norm=True,
                 stochastic_mode=False, scale_attention=True, triangular_masking=True, local_attention=False,
                 window_size=256, return_tuple=True, moe_experts=1, global_experts=1, k=1, capacity_factor=1.,
                 eval_capacity_factor=1., min_capacity=1, noisy_gate_policy=None, drop_tokens=True, use_rts=False,
                 mlp_type="standard", scale_attn_by_inverse_layer_idx=False):
        dtype = torch.int8 if q_int8 else (torch.float16 if fp16 else (torch.bfloat16 if bf16 else torch.float32))
        super().__init__(hidden_size, intermediate_size, heads, num_hidden_layers, layer_norm_eps, local_rank, mp_size,
                         dtype, pre_layer_norm, stochastic_mode=stochastic_mode, scale_attention=scale_attention,
                         triangular_masking=triangular_masking, local_attention=local_attention,
                         window_size=window_size, return_tuple=return_tuple,
                         scale_attn_by_inverse_layer_idx=scale_attn_by_inverse_layer_idx)
        self.moe_experts, self.global_experts, self.k = moe_experts, global_experts, k
        self.capacity_factor, self.eval_capacity_factor, self.min_capacity = capacity_factor, eval_capacity_factor, min_capacity
        self.noisy_gate_policy, self.drop_tokens, self.use_rts, self.mlp_type = noisy_gate_policy, drop_tokens, use_rts, mlp_type


class _ExpertMLP(nn.Module):

    def __init__(self, cfg):
        super().__init__()
        h, inter = cfg.hidden_size, cfg.intermediate_size // cfg.mp_size
        self.inter_w, self.inter_b = _empty(cfg, inter, h), _empty(cfg, inter)
        self.output_w, self.output_b = _empty(cfg, h, inter), _empty(cfg, h)

    def forward(self, x):
        return F.linear(F.gelu(F.linear(x, self.inter_w, self.inter_b), approximate="tanh"), self.output_w, self.output_b)


class DeepSpeedMoEInference(nn.Module):
    layer_id = 0

    def __init__(self, config, mp_group=None, ep_group=None, expert_mp_group=None, quantize_scales=None,
                 quantize_groups=1, merge_count=1, mlp_extra_grouping=False):
        super().__init__()
        self.config = config
        self.config.layer_id = DeepSpeedMoEInference.layer_id
        DeepSpeedMoEInference.layer_id += 1
        self.attention = DeepSpeedSelfAttention(config, mp_group)
        self.attn_nw, self.attn_nb = _empty(config, config.hidden_size), _empty(config, config.hidden_size)
        self.norm_w, self.norm_b = _empty(config, config.hidden_size), _empty(config, config.hidden_size)
        self.ep_group, self.mp_group = ep_group, mp_group
        self.num_local = max(1, config.moe_experts)
        self.experts = nn.ModuleList([_ExpertMLP(config) for _ in range(self.num_local)])
        self.gate_w = _empty(config, config.global_experts, config.hidden_size)
        if config.mlp_type == "residual":
            self.res_mlp = _ExpertMLP(config)
            self.res_coef = _empty(config, 2, config.hidden_size)

    def _moe(self, x):
        """Dropless top-k over the global experts; with an expert-parallel group each rank computes the experts it owns
        for all tokens of the group (all-gather tokens, all-reduce the weighted outputs)."""
        cfg = self.config
        shape = x.shape
        tok = x.reshape(-1, shape[-1])
        ep = dist.get_world_size(self.ep_group) if (self.ep_group is not None and dist.is_initialized()) else 1
        rank = dist.get_rank(self.ep_group) if ep > 1 else 0
        if ep > 1:
            parts = [torch.empty_like(tok) for _ in range(ep)]
            dist.all_gather(parts, tok.contiguous(), group=self.ep_group)
            tok_all = torch.cat(parts, dim=0)
        else:
            tok_all = tok
        probs = F.softmax(F.linear(tok_all.float(), self.gate_w.float()), dim=-1)
        w, idx = probs.topk(cfg.k, dim=-1)
        out = torch.zeros_like(tok_all, dtype=torch.float32)
        for local, expert in enumerate(self.experts):
            gid = rank * self.num_local + local
            hit = (idx == gid)
            rows = hit.any(dim=-1).nonzero(as_tuple=True)[0]
            if rows.numel():
                wt = (w * hit)[rows].sum(dim=-1, keepdim=True)
                out[rows] += wt * expert(tok_all[rows]).float()
        if ep > 1:
            dist.all_reduce(out, group=self.ep_group)
            n = tok.shape[0]
            out = out[rank * n:(rank + 1) * n]
        return out.to(x.dtype).reshape(shape)

    def forward(self, input, input_mask=None, attention_mask=None, head_mask=None, layer_past=None,
                get_key_value=False, get_present=False, encoder_output=None, enc_dec_attn_mask=None,
                encoder_hidden_states=None, encoder_attention_mask=None, use_cache=False, output_attentions=False):
        cfg = self.config
        get_present = get_present or get_key_value or use_cache
        mask = input_mask if attention_mask is None else attention_mask
        h = input.to(self.attention.attn_qkvw.dtype)
        a_in = _norm(cfg, h, self.norm_w, self.norm_b)
        attn_out, present = self.attention(a_in, mask, layer_past, get_present)
        h = h + attn_out + self.attention.attn_ob
        m_in = _norm(cfg, h, self.attn_nw, self.attn_nb)
        moe_out = self._moe(m_in)
        if cfg.mlp_type == "residual":
            coef = F.softmax(F.linear(m_in.float(), self.res_coef.float()), dim=-1).to(m_in.dtype)
            moe_out = moe_out * coef[..., 0:1] + self.res_mlp(m_in) * coef[..., 1:2]
        out = (h + moe_out).to(input.dtype)
        if get_present:
            return (out, present)
        return (out, ) if cfg.return_tuple else out
