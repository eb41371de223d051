"""Post-training int8 weight quantization for inference checkpoints (reference ``runtime/weight_quantizer.py:10``):
group-wise symmetric quantization of attention/MLP weight matrices while a checkpoint is being loaded/merged."""
import torch


class WeightQuantization:

    def __init__(self, mlp_extra_grouping=True, mp_size=1):
        self.dense_scales, self.qkv_scales, self.mlp4hh_scales, self.mlph4h_scales = [], [], [], []
        self.mlp_extra_grouping = mlp_extra_grouping
        self.mp_size = mp_size

    def quantize_data(self, data, quantize_bits, groups, key=None):
        groups = max(1, groups)
        parts = torch.split(data.float().flatten(), data.numel() // groups)
        max_d = [max(float(g.max()), float(g.min().abs())) for g in parts]
        scales = [(float(1 << quantize_bits) / (2 * mx + 1e-5)) for mx in max_d]
        q = [(g * s).round().clamp(-(1 << (quantize_bits - 1)), (1 << (quantize_bits - 1)) - 1)
             for g, s in zip(parts, scales)]
        return torch.cat(q).reshape(data.shape).to(torch.int8), torch.tensor([[1.0 / s for s in scales]])

    def is_mlp(self, data, merge_count=1):
        return (self.mp_size * data.shape[0] * merge_count) / data.shape[1] == 4 or \
            (self.mp_size * data.shape[1] * merge_count) / data.shape[0] == 4

    def is_qkv(self, data):
        return (self.mp_size * data.shape[0]) / data.shape[1] == 3 or \
            (self.mp_size * data.shape[1]) / data.shape[0] == 3

    def Quantize(self, value_list, quantize_bits, groups, key, merge_dim=0):
        if self.mlp_extra_grouping and self.is_mlp(value_list[0], merge_count=len(value_list)):
            groups *= 2
        q_scale = []
        for index, data in enumerate(value_list):
            data_int, data_scale = self.quantize_data(data, quantize_bits, groups, key)
            q_scale.append(data_scale)
            value_list[index] = data_int
        q_scale = 1 / torch.cat(q_scale, dim=merge_dim).view(-1).unsqueeze(0)
        if "mlp.dense_4h_to_h.weight" in key:
            self.mlp4hh_scales.append(q_scale)
        elif "mlp.dense_h_to_4h.weight" in key:
            self.mlph4h_scales.append(q_scale)
        elif "attention.query_key_value.weight" in key:
            self.qkv_scales.append(q_scale)
        else:
            self.dense_scales.append(q_scale)
        return value_list

    def merge_layer_scales(self, layer_scales):
        max_dim = max(s.numel() for s in layer_scales)
        padded = [torch.cat((s, torch.zeros(1, max_dim - s.numel())), dim=-1) if s.numel() < max_dim else s
                  for s in layer_scales]
        return torch.cat(padded).unsqueeze(0)

    def merge_scales(self):
        return torch.cat([self.merge_layer_scales([qkv, dense, h4h, fhh]) for dense, qkv, fhh, h4h in
                          zip(self.dense_scales, self.qkv_scales, self.mlp4hh_scales, self.mlph4h_scales)])

    def merge_scales_split(self, split_count):
        out = [[] for _ in range(split_count)]
        for dense, qkv, fhh, h4h in zip(self.dense_scales, self.qkv_scales, self.mlp4hh_scales, self.mlph4h_scales):
            dense_s = torch.split(dense, dense.numel() // split_count)
            qkv_s = torch.split(qkv, qkv.numel() // split_count)
            fhh_s = torch.split(fhh, fhh.numel() // split_count)
            h4h_s = torch.split(h4h, h4h.numel() // split_count)
            for s in range(split_count):
                out[s].append(torch.cat([torch.cat((qkv_s[s], torch.zeros_like(qkv_s[s])), dim=1),
                                         torch.cat((dense_s[s], torch.zeros_like(dense_s[s])), dim=1),
                                         h4h_s[s], fhh_s[s]]).unsqueeze(0))
            for scales in out:
                scales[-1] = scales[-1]
        return [torch.cat(s) for s in out]

    def sd_quantize_megatron(self, sd, quantize_bits, groups):
        keys = sd.keys()
        for key in keys:
            value_list = [sd[key]]
            if "attention.dense.weight" in key or "mlp.dense_4h_to_h.weight" in key or \
                    "mlp.dense_h_to_4h.weight" in key or "attention.query_key_value.weight" in key:
                value_list = self.Quantize(value_list, quantize_bits, groups, key=key)
            sd[key] = value_list[0]
        return sd, self.merge_scales()

    def model_quantize(self, model, quantize_policy, quantize_bits, groups):
        all_scales = []

        def quantize_fn(layer, policy_cls):
            policy = policy_cls(layer)
            _, qkvw, _, dense_w, _, _ = policy.attention()
            _, _h4h_w, _, _4hh_w, _ = policy.mlp()
            layer_scales = []
            for key, w in (("attention.query_key_value.weight", qkvw), ("attention.dense.weight", dense_w),
                           ("mlp.dense_h_to_4h.weight", _h4h_w), ("mlp.dense_4h_to_h.weight", _4hh_w)):
                if w is not None:
                    data_q, scale = self.quantize_data(w.data, quantize_bits, groups, key)
                    w.data.copy_(data_q.to(w.dtype))
                    layer_scales.append(1 / scale.to(w.device).view(-1).unsqueeze(0))
            all_scales.append(self.merge_layer_scales(layer_scales))
            return layer

        def _walk(module):
            for name, child in module.named_children():
                if child.__class__ in quantize_policy:
                    setattr(module, name, quantize_fn(child, quantize_policy[child.__class__]))
                else:
                    _walk(child)
            return module

        return _walk(model), torch.cat(all_scales) if all_scales else torch.zeros(0)
