"""Engine construction for FastGen (reference ``inference/v2/engine_factory.py``, ``config_v2.py``,
``checkpoint/huggingface_engine.py``): from a local HuggingFace model directory or from a checkpoint written by
``InferenceEngineV2.serialize``."""
import glob
import json
import logging
import os
import pickle

import torch

from ...config.base import DeepSpeedConfigModel
from .engine_v2 import InferenceEngineV2, RaggedInferenceEngineConfig  # noqa: F401
from .model_implementations.policies import build_hf_model

DS_MODEL_CONFIG = "ds_model_config.pkl"


class DeepSpeedTPConfig(DeepSpeedConfigModel):
    tp_size: int = 1


def make_param_filename(base, tp_rank, tp_size):
    return os.path.join(base, f"params_rank_{tp_rank}_of_{tp_size}.pt")


def make_metadata_filename(base, tp_rank, tp_size):
    return os.path.join(base, f"metadata_rank_{tp_rank}_of_{tp_size}.json")


def _load_hf_state_dict(path):
    """All tensors of a local HF checkpoint directory (safetensors shards preferred, then .bin)."""
    sd = {}
    shards = sorted(glob.glob(os.path.join(path, "*.safetensors")))
    if shards:
        from safetensors.torch import load_file
        for f in shards:
            sd.update(load_file(f))
        return sd
    shards = sorted(glob.glob(os.path.join(path, "pytorch_model*.bin")))
    if not shards:
        raise FileNotFoundError(f"no *.safetensors or pytorch_model*.bin under {path}")
    for f in shards:
        sd.update(torch.load(f, map_location="cpu", weights_only=True))
    return sd


def _device_dtype(engine_config):
    dev = torch.device("cuda", torch.cuda.current_device()) if torch.cuda.is_available() else torch.device("cpu")
    dtype = getattr(engine_config, "dtype", None) or (torch.bfloat16 if dev.type == "cuda" else torch.float32)
    return dev, dtype


def build_engine_from_ds_checkpoint(path, engine_config=None, debug_level=logging.INFO):
    """Engine from the directory ``InferenceEngineV2.serialize`` wrote (model config pickle + parameter file)."""
    engine_config = engine_config or RaggedInferenceEngineConfig()
    tp = getattr(getattr(engine_config, "tensor_parallel", None), "tp_size", 1)
    meta = json.load(open(make_metadata_filename(path, 0, tp)))
    with open(os.path.join(path, DS_MODEL_CONFIG), "rb") as f:
        hf_config = pickle.load(f)
    if meta.get("policy") != hf_config.model_type:
        raise ValueError(f"Unknown policy {meta.get('policy')} for model {path}")
    dev, dtype = _device_dtype(engine_config)
    sd = torch.load(make_param_filename(path, 0, tp), map_location="cpu", weights_only=True)
    model = build_hf_model(hf_config, None, dtype=dtype, device=dev)
    model.load_state_dict(sd)
    model.hf_config = hf_config
    return InferenceEngineV2(model, engine_config)


def build_hf_engine(path, engine_config=None, debug_level=logging.INFO):
    """Engine for a HuggingFace causal LM stored under ``path`` (a model directory with ``config.json`` and weight
    shards; hub names need network access and are not resolved here), or for a serialized FastGen checkpoint."""
    engine_config = engine_config or RaggedInferenceEngineConfig()
    if os.path.exists(os.path.join(path, DS_MODEL_CONFIG)):
        return build_engine_from_ds_checkpoint(path, engine_config, debug_level)
    if not os.path.isdir(path):
        raise FileNotFoundError(f"{path} is not a local model directory")
    from transformers import AutoConfig
    hf_config = AutoConfig.from_pretrained(path, local_files_only=True)
    if hf_config.model_type == "opt" and not getattr(hf_config, "do_layer_norm_before", True):
        raise ValueError("Detected OPT-350m model. This model is not currently supported.")
    dev, dtype = _device_dtype(engine_config)
    model = build_hf_model(hf_config, _load_hf_state_dict(path), dtype=dtype, device=dev)
    model.hf_config = hf_config
    return InferenceEngineV2(model, engine_config)
