"""Load Megatron-style model-parallel checkpoints at a different inference TP degree (reference
``runtime/state_dict_factory.py:22``): merge ``mp`` shards when the checkpoint has more ranks than the run, split them
when it has fewer; QKV tensors honour the checkpoint version's layout; optional int8 quantization on the fly."""
import copy
import json
import os
from abc import ABC, abstractmethod

import torch

from ..utils.logging import logger
from .weight_quantizer import WeightQuantization

AUTO_MODULE_KEY = "auto"


class SDLoaderFactory:

    @staticmethod
    def get_sd_loader_json(json_file, checkpoint_engine=None):
        data = json_file
        if isinstance(json_file, str):
            with open(json_file) as f:
                data = json.load(f)
        sd_type = data["type"]
        ckpt_list = data["checkpoints"]
        version = data["version"]
        mp_size = data.get("mp_size", 0)
        if sd_type.lower() in ("bloom", "ds_model"):
            return data
        return SDLoaderFactory.get_sd_loader(ckpt_list, checkpoint_engine, sd_type, version), mp_size

    @staticmethod
    def get_sd_loader(ckpt_list, checkpoint_engine=None, sd_type="Megatron", version=None):
        if sd_type == "Megatron":
            return MegatronSDLoader(ckpt_list, version, checkpoint_engine)
        raise ValueError(f"{sd_type} checkpoint type is not supported")


class SDLoaderBase(ABC):

    def __init__(self, ckpt_list, version, checkpoint_engine=None):
        self.module_key = None
        self.ckpt_list = ckpt_list
        self.version = version
        self.checkpoint_engine = checkpoint_engine
        self.check_ckpt_list()

    def _load_file(self, path):
        if self.checkpoint_engine is not None:
            return self.checkpoint_engine.load(path, map_location=lambda storage, loc: storage)
        return torch.load(path, map_location="cpu", weights_only=False)

    def load(self, mp_world_size, mp_rank, module_key=AUTO_MODULE_KEY, is_pipe_parallel=False, quantize=False,
             quantize_bits=8, quantize_groups=64, mlp_extra_grouping=True):
        self.module_key = module_key
        num_ckpt = len(self.ckpt_list)
        idx = mp_rank * num_ckpt // mp_world_size
        if is_pipe_parallel and module_key is not None and mp_world_size != num_ckpt:
            mp_world_size = num_ckpt
            idx = 0
        load_path = self.ckpt_list[idx]
        merge_count = 1
        if num_ckpt == mp_world_size:
            assert os.path.exists(load_path)
            sd = self._load_file(load_path)
            if quantize:
                quantizer = WeightQuantization(mlp_extra_grouping=mlp_extra_grouping, mp_size=mp_world_size)
                sd_module, all_scales = quantizer.sd_quantize_megatron(self.get_module(sd), quantize_bits,
                                                                       quantize_groups)
                self.set_module(sd, sd_module)
            else:
                all_scales = None
        elif num_ckpt > mp_world_size:
            sd, all_scales, merge_count = self.merge_state_dict(mp_world_size, mp_rank, quantize, quantize_bits,
                                                                quantize_groups, mlp_extra_grouping)
        else:
            sd, all_scales = self.split_state_dict(mp_world_size, mp_rank, quantize, quantize_bits, quantize_groups,
                                                   mlp_extra_grouping)
        return load_path, sd, (all_scales, merge_count)

    def get_merge_state_dicts(self, mp_world_size, mp_rank):
        num_ckpt = len(self.ckpt_list)
        assert num_ckpt % mp_world_size == 0, "Invalid checkpoints and world size for sd merge"
        num_to_merge = num_ckpt // mp_world_size
        ckpts = self.ckpt_list[mp_rank * num_to_merge:(mp_rank + 1) * num_to_merge]
        logger.info(f"mp_rank: {mp_rank}, ckpt_list: {ckpts}")
        return [self._load_file(c) for c in ckpts]

    def get_split_state_dict(self, mp_world_size, mp_rank):
        num_ckpt = len(self.ckpt_list)
        assert mp_world_size % num_ckpt == 0, "Invalid checkpoints and world size for sd split"
        num_to_split = mp_world_size // num_ckpt
        ckpt_index = mp_rank // num_to_split
        ckpt_offset = mp_rank % num_to_split
        return self._load_file(self.ckpt_list[ckpt_index]), num_to_split, ckpt_offset

    def _choose_module_key(self, sd):
        assert not ("module" in sd and "model" in sd), "checkpoint has both 'model' and 'module' keys"
        assert "module" in sd or "model" in sd, "checkpoint contains neither 'model' or 'module' keys"
        return "module" if "module" in sd else "model"

    def get_module(self, sd):
        if self.module_key is None:
            return sd
        if self.module_key == AUTO_MODULE_KEY:
            return sd[self._choose_module_key(sd)]
        return sd[self.module_key]

    def set_module(self, sd, module):
        if self.module_key is None:
            sd = module
        elif self.module_key == AUTO_MODULE_KEY:
            sd[self._choose_module_key(sd)] = module
        else:
            sd[self.module_key] = module
        return sd

    def check_ckpt_list(self):
        assert len(self.ckpt_list) > 0
        sd = self._load_file(self.ckpt_list[0])
        if "mp_world_size" in sd:
            assert len(self.ckpt_list) == sd["mp_world_size"], \
                f"checkpoint count {len(self.ckpt_list)} is different from saved mp_world_size {sd['mp_world_size']}"

    @abstractmethod
    def merge_state_dict(self, mp_world_size, mp_rank, quantize, quantize_bits, groups, mlp_extra_grouping):
        ...

    @abstractmethod
    def split_state_dict(self, mp_world_size, mp_rank, quantize, quantize_bits, groups, mlp_extra_grouping):
        ...

    @abstractmethod
    def sanity_check(self, ckpt_file_name):
        ...


class MegatronSDLoader(SDLoaderBase):
    """Column-parallel tensors (QKV, h_to_4h, embeddings) concatenate on dim 0, row-parallel ones (attention dense,
    4h_to_h) on dim 1; fused QKV is re-interleaved according to the checkpoint version (0: [q|k|v] per shard,
    1/2: per-head interleaved -> plain concatenation)."""

    def merge_query_key_value(self, param_list, ckpt_ver):
        if ckpt_ver == 0:
            assert param_list[0].shape[0] % 3 == 0
            size_qkv = param_list[0].shape[0] // 3
            split = [torch.split(p, size_qkv, dim=0) for p in param_list]
            return torch.cat([torch.cat([s[i] for s in split], dim=0) for i in range(3)], dim=0)
        if ckpt_ver in (1.0, 2.0):
            return torch.cat(param_list, dim=0)
        raise AssertionError(f"checkpoint version: {ckpt_ver} is not supported")

    def split_query_key_value(self, param, num_to_split, offset, ckpt_ver):
        if ckpt_ver == 0:
            assert param.shape[0] % 3 == 0
            size_qkv = param.shape[0] // 3
            parts = torch.split(param, size_qkv, dim=0)
            assert parts[0].shape[0] % num_to_split == 0
            sz = parts[0].shape[0] // num_to_split
            return torch.cat([torch.split(p, sz, dim=0)[offset] for p in parts], dim=0)
        if ckpt_ver in (1.0, 2.0):
            assert param.shape[0] % num_to_split == 0
            return torch.split(param, param.shape[0] // num_to_split, dim=0)[offset]
        raise AssertionError(f"checkpoint version: {ckpt_ver} is not supported")

    def merge_state_dict(self, mp_world_size, mp_rank, quantize=False, quantize_bits=8, groups=64,
                         mlp_extra_grouping=True):
        self.sanity_check(self.ckpt_list[0])
        sd_list = self.get_merge_state_dicts(mp_world_size, mp_rank)
        ds_sd = copy.deepcopy(sd_list[0])
        client = [self.get_module(sd) for sd in sd_list]
        ckpt_ver = self.get_checkpoint_version(ds_sd)
        quantizer = WeightQuantization(mlp_extra_grouping, mp_world_size) if quantize else None
        new_sd = type(client[0])()
        for key in client[0].keys():
            values = [sd[key] for sd in client]
            if "attention.dense.weight" in key or "mlp.dense_4h_to_h.weight" in key:
                if quantize:
                    values = quantizer.Quantize(values, quantize_bits, groups, key=key, merge_dim=1)
                new_sd[key] = torch.cat(values, dim=1)
            elif "attention.query_key_value" in key:
                if quantize and "attention.query_key_value.weight" in key:
                    values = quantizer.Quantize(values, quantize_bits, groups, key=key)
                    new_sd[key] = torch.cat(values, dim=0)
                else:
                    new_sd[key] = self.merge_query_key_value(values, ckpt_ver)
            elif "mlp.dense_h_to_4h.weight" in key or "word_embeddings.weight" in key or \
                    "mlp.dense_h_to_4h.bias" in key:
                if quantize and "mlp.dense_h_to_4h.weight" in key:
                    values = quantizer.Quantize(values, quantize_bits, groups, key=key)
                new_sd[key] = torch.cat(values, dim=0)
            else:
                new_sd[key] = values[0]
        all_scales = quantizer.merge_scales() if quantize else None
        return self.set_module(ds_sd, new_sd), all_scales, len(client)

    def split_state_dict(self, mp_world_size, mp_rank, quantize=False, quantize_bits=8, groups=64,
                         mlp_extra_grouping=True):
        sd, num_to_split, offset = self.get_split_state_dict(mp_world_size, mp_rank)
        ds_sd = copy.deepcopy(sd)
        client = self.get_module(sd)
        ckpt_ver = self.get_checkpoint_version(ds_sd)
        quantizer = WeightQuantization(mlp_extra_grouping, mp_world_size) if quantize else None
        new_sd = type(client)()
        for key, value in client.items():
            if "attention.dense.weight" in key or "mlp.dense_4h_to_h.weight" in key:
                assert value.shape[1] % num_to_split == 0
                v = torch.split(value, value.shape[1] // num_to_split, dim=1)[offset].contiguous()
                new_sd[key] = quantizer.Quantize([v], quantize_bits, groups, key)[0] if quantize else v
            elif "attention.query_key_value" in key:
                v = self.split_query_key_value(value, num_to_split, offset, ckpt_ver)
                if quantize and "attention.query_key_value.weight" in key:
                    v = quantizer.Quantize([v], quantize_bits, groups, key)[0]
                new_sd[key] = v
            elif "mlp.dense_h_to_4h.weight" in key or "word_embeddings.weight" in key or \
                    "mlp.dense_h_to_4h.bias" in key or "final_linear.weight" in key:
                assert value.shape[0] % num_to_split == 0
                v = torch.split(value, value.shape[0] // num_to_split, dim=0)[offset].contiguous()
                if quantize and "mlp.dense_h_to_4h.weight" in key:
                    v = quantizer.Quantize([v], quantize_bits, groups, key)[0]
                new_sd[key] = v
            else:
                new_sd[key] = value
        all_scales = quantizer.merge_scales_split(num_to_split) if quantize else None
        return self.set_module(ds_sd, new_sd), all_scales

    def sanity_check(self, ckpt_file_name):
        sd = self.get_module(self._load_file(ckpt_file_name))
        for partial in ("attention.dense.weight", "mlp.dense_4h_to_h.weight", "attention.query_key_value",
                        "mlp.dense_h_to_4h.weight", "mlp.dense_h_to_4h.bias"):
            assert any(partial in k for k in sd.keys()), \
                f'key: {partial} is not found in the checkpoint {ckpt_file_name}'

    def get_checkpoint_version(self, state_dict):
        return self.version if self.version is not None else state_dict.get("checkpoint_version", 0)
