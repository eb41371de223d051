"""Module surgery for compression (reference ``compression/helper.py``): replace Linear/Conv2d/Embedding/BatchNorm by
their ``*_Compress`` twins, enable techniques per config group, bake masks in with ``fix_compression``."""
import torch

from .basic_layer import (BNLayer_Compress, ColumnParallelLinear_Compress, Conv2dLayer_Compress, Embedding_Compress,
                          LinearLayer_Compress, RowParallelLinear_Compress)
from .constants import *  # noqa: F401,F403


def recursive_getattr(model, module_name):
    for part in module_name.split("."):
        model = getattr(model, part)
    return model


def recursive_setattr(model, module_name, module):
    parts = module_name.split(".")
    for part in parts[:-1]:
        model = getattr(model, part)
    setattr(model, parts[-1], module)


def _twin_of(old, mpu):
    if isinstance(old, (LinearLayer_Compress, Conv2dLayer_Compress, Embedding_Compress, BNLayer_Compress)):
        return old
    if mpu is not None and isinstance(old, mpu.ColumnParallelLinear):
        new = ColumnParallelLinear_Compress(mpu, old.input_size, old.output_size, gather_output=old.gather_output,
                                            skip_bias_add=old.skip_bias_add, bias=old.bias is not None)
    elif mpu is not None and isinstance(old, mpu.RowParallelLinear):
        new = RowParallelLinear_Compress(mpu, old.input_size, old.output_size, input_is_parallel=old.input_is_parallel,
                                         skip_bias_add=old.skip_bias_add, bias=old.bias is not None)
    elif isinstance(old, torch.nn.Linear):
        new = LinearLayer_Compress(old.in_features, old.out_features, bias=old.bias is not None)
    elif isinstance(old, torch.nn.Conv2d):
        new = Conv2dLayer_Compress(old.in_channels, old.out_channels, old.kernel_size, old.stride, old.padding,
                                   old.dilation, old.groups, old.bias is not None, old.padding_mode)
    elif isinstance(old, torch.nn.BatchNorm2d):
        new = BNLayer_Compress(old.num_features, old.eps, old.momentum, old.affine, old.track_running_stats)
        new = new.to(old.weight.device, old.weight.dtype)
        new.load_state_dict(old.state_dict())
        return new
    elif isinstance(old, torch.nn.Embedding):
        new = Embedding_Compress(old.num_embeddings, old.embedding_dim, old.padding_idx, old.max_norm, old.norm_type,
                                 old.scale_grad_by_freq, old.sparse)
    else:
        return None
    new = new.to(old.weight.device, old.weight.dtype)
    new.weight.data = old.weight.data
    if getattr(old, "bias", None) is not None:
        new.bias.data = old.bias.data
    return new


def module_replacement(model, module_name, compression_technique=None, mpu=None):
    """Swap ``module_name`` for its compression-aware twin and enable the techniques given as
    ``{technique: params}``."""
    old = recursive_getattr(model, module_name)
    new = _twin_of(old, mpu)
    if new is None:
        return None
    for technique, p in (compression_technique or {}).items():
        if technique == SPARSE_PRUNING:
            if p[SPARSE_PRUNING_ENABLED]:
                new.enable_sparse_pruning(p[SPARSE_PRUNING_DENSE_RATIO], p[SPARSE_PRUNING_METHOD])
        elif technique == ROW_PRUNING:
            if p[ROW_PRUNING_ENABLED]:
                new.enable_row_pruning(p[ROW_PRUNING_DENSE_RATIO], p[ROW_PRUNING_METHOD])
        elif technique == HEAD_PRUNING:
            if p[HEAD_PRUNING_ENABLED]:
                new.enable_head_pruning(p[HEAD_PRUNING_DENSE_RATIO], p[HEAD_PRUNING_METHOD],
                                        p[HEAD_PRUNING_NUM_HEADS])
        elif technique == CHANNEL_PRUNING:
            if p[CHANNEL_PRUNING_ENABLED]:
                new.enable_channel_pruning(p[CHANNEL_PRUNING_DENSE_RATIO], p[CHANNEL_PRUNING_METHOD])
        elif technique == ACTIVATION_QUANTIZATION:
            if p[ACTIVATION_QUANTIZATION_ENABLED]:
                new.enable_activation_quantization(p[ACTIVATION_QUANTIZE_BITS], p[ACTIVATION_QUANTIZE_TYPE],
                                                   p[ACTIVATION_QUANTIZE_RANGE])
        elif technique == WEIGHT_QUANTIZATION:
            if p[WEIGHT_QUANTIZE_ENABLED]:
                new.enable_weight_quantization(p[WEIGHT_QUANTIZE_START_BITS], p[WEIGHT_QUANTIZE_TARGET_BITS],
                                               p[WEIGHT_QUANTIZATION_PERIOD], p[WEIGHT_QUANTIZE_IN_FORWARD_ENABLED],
                                               p[WEIGHT_QUANTIZE_TYPE], p[WEIGHT_QUANTIZE_GROUPS])
        else:
            raise NotImplementedError(f"Compression technique {technique} is not implemented")
    recursive_setattr(model, module_name, new)
    return new


def is_module_compressible(module, mpu=None):
    ok = isinstance(module, (torch.nn.Linear, torch.nn.Conv2d, torch.nn.Embedding, torch.nn.BatchNorm2d))
    if mpu is not None:
        ok = ok or isinstance(module, (mpu.RowParallelLinear, mpu.ColumnParallelLinear))
    return ok


def compression_preparation(model, compression_technique_list, mpu):
    """First give every compressible module its twin (no technique), then enable the configured techniques."""
    for name, module in list(model.named_modules()):
        if name and is_module_compressible(module, mpu):
            module_replacement(model, name, mpu=mpu)
    for module_name_lists, _, technique in compression_technique_list:
        for names in module_name_lists:
            for name in names:
                module_replacement(model, name, technique)
    return model


def fix_compression(model, module_name, compression_technique, mask=None, dim_reduction=False):
    """Make the technique permanent on one module; returns the mask for related (downstream) modules."""
    module = recursive_getattr(model, module_name)
    for technique, p in compression_technique.items():
        if technique == WEIGHT_QUANTIZATION and p[WEIGHT_QUANTIZE_IN_FORWARD_ENABLED] and p[WEIGHT_QUANTIZE_ENABLED]:
            return module.fix_weight_quantization()
        if technique == SPARSE_PRUNING and p[SPARSE_PRUNING_ENABLED]:
            return module.fix_sparse_pruning_helper()
        if technique == ROW_PRUNING and (p[ROW_PRUNING_ENABLED] or mask is not None):
            return module.fix_row_col_pruning_helper(mask, dim_reduction=dim_reduction)
        if technique == HEAD_PRUNING and (p[HEAD_PRUNING_ENABLED] or mask is not None):
            return module.fix_head_pruning_helper(mask, p[HEAD_PRUNING_NUM_HEADS], dim_reduction=dim_reduction)
        if technique == CHANNEL_PRUNING and (p[CHANNEL_PRUNING_ENABLED] or mask is not None):
            return module.fix_channel_pruning_helper(mask, dim_reduction=dim_reduction)
    return None


def convert_conv1d_to_linear(model, convert_type):
    """HF GPT-2 ``Conv1D`` (weight [in, out]) -> ``nn.Linear`` so that the linear techniques apply."""
    c_model = model.module if hasattr(model, "module") else model
    for name, module in list(c_model.named_modules()):
        if isinstance(module, convert_type):
            old = recursive_getattr(c_model, name)
            new = torch.nn.Linear(old.weight.data.size(0), old.weight.data.size(1), bias=old.bias is not None)
            new.weight.data = old.weight.data.t().contiguous()
            if new.bias is not None:
                new.bias.data = old.bias.data.view(-1)
            recursive_setattr(c_model, name, new)
    return model


def generate_pruners(config, model):
    """snip_momentum block pruning is delegated to Intel neural_compressor in the reference; same here."""
    try:
        from neural_compressor.compression.pruner import pruners as nc_pruners
        from neural_compressor.compression.pruner.utils import parse_to_prune, process_config
    except ImportError as e:
        raise ImportError("snip_momentum sparse pruning needs the neural_compressor package") from e
    out = []
    for info in process_config(config):
        modules = parse_to_prune(info, model)
        if modules:
            out.append(nc_pruners.get_pruner(info, modules))
            info["modules"] = list(modules)
            info["len_of_modules"] = len(info["modules"])
    return out


def register_on_step_begin(model):

    def hook(module, input):
        for pruner in module.pruners:
            pruner.on_step_begin(0)

    return model.register_forward_pre_hook(hook)


def rewrite_optimizer_step(opt):

    def new_step(self, closure=None):
        if hasattr(self, "pruners"):
            for pruner in self.pruners:
                pruner.on_before_optimizer_step()
        res = self.orig_step(closure) if closure is not None else self.orig_step()
        if hasattr(self, "pruners"):
            for pruner in self.pruners:
                pruner.on_after_optimizer_step()
        return res

    import types
    opt.orig_step = opt.step
    opt.step = types.MethodType(new_step, opt)
    return opt
