"""Model surgery for random-LTD (reference ``runtime/data_pipeline/data_routing/helper.py``)."""
from collections import OrderedDict

from .basic_layer import RandomLayerTokenDrop


def _get(model, name):
    for part in name.split("."):
        model = getattr(model, part)
    return model


def _set(model, name, module):
    parts = name.split(".")
    for part in parts[:-1]:
        model = getattr(model, part)
    setattr(model, parts[-1], module)


def convert_to_random_ltd(model, convert_type):
    """Wrap every submodule of ``convert_type`` of ``model`` (an engine or a module) in ``RandomLayerTokenDrop``."""
    c_model = model.module if hasattr(model, "module") else model
    for name, module in list(c_model.named_modules()):
        if isinstance(module, convert_type) and name:
            _set(c_model, name, RandomLayerTokenDrop(_get(c_model, name)))
    if hasattr(model, "random_ltd_initialize"):
        model.random_ltd_initialize()
    return model


def remove_random_ltd_state_dict(state_dict):
    return OrderedDict((k.replace(".random_ltd_layer", ""), v) for k, v in state_dict.items())


def save_without_random_ltd(model):
    c_model = model.module if hasattr(model, "module") else model
    return remove_random_ltd_state_dict(c_model.state_dict())
