"""ZeRO-3 leaf modules (reference ``utils/z3_leaf_module.py``): a flagged module and everything below it form ONE
flat unit, gathered when the module is entered - for blocks whose children run in data-dependent order (MoE experts)
or are too fine-grained for per-module hooks. The flag is the instance attribute ``_z3_leaf`` that
``runtime/zero/units.py`` reads when it cuts the model into units."""
from typing import List, Optional, Tuple, Type, Union

import torch

from .logging import logger


def z3_leaf_module(model: torch.nn.Module) -> bool:
    return bool(getattr(model, "_z3_leaf", False))


def z3_leaf_parameter(model: torch.nn.Parameter) -> bool:
    """True for a parameter that lives inside a leaf module (set when the units are built)."""
    return bool(getattr(model, "ds_z3_leaf_module", None) is not None)


def get_z3_leaf_modules(model: torch.nn.Module) -> List[torch.nn.Module]:
    return [m for m in model.modules() if z3_leaf_module(m)]


def set_z3_leaf_module(model: torch.nn.Module, flag: bool):
    model._z3_leaf = flag


def _class_matches(module, spec):
    if isinstance(spec, type):
        return isinstance(module, spec)
    cls = type(module)
    return spec == cls.__name__ or spec == f"{cls.__module__}.{cls.__qualname__}"


def _flag_by_class(model, specs, flag, raise_if_not_found):
    specs = list(specs)
    bad = [s for s in specs if not isinstance(s, (type, str))]
    assert not bad, f"leaf_module_classes must be a list of types or names, got {specs}"
    hit = []
    for m in model.modules():
        if any(_class_matches(m, s) for s in specs):
            m._z3_leaf = flag
            hit.append(m)
    if not hit and raise_if_not_found:
        raise ValueError(f"No modules of type {specs} found in model {model}")
    return hit


def set_z3_leaf_modules(model: torch.nn.Module, leaf_module_classes: Union[List[Type], List[str]],
                        raise_if_not_found: bool = True) -> List[torch.nn.Module]:
    """Flag every module whose class (type, class name or fully qualified name) is listed."""
    return _flag_by_class(model, leaf_module_classes, True, raise_if_not_found)


def unset_z3_leaf_modules(model: torch.nn.Module, leaf_module_classes: Union[List[Type], List[str]],
                          raise_if_not_found: bool = True) -> List[torch.nn.Module]:
    return _flag_by_class(model, leaf_module_classes, False, raise_if_not_found)


def set_z3_leaf_modules_by_name(model: torch.nn.Module, module_names: List[str], flag: bool = True,
                                raise_if_not_found: bool = True) -> Tuple[List[torch.nn.Module], List[str]]:
    """Flag modules by their ``named_modules()`` names. Returns (matched modules, names that matched nothing)."""
    table = dict(model.named_modules())
    hit, missing = [], []
    for name in module_names:
        if name in table:
            table[name]._z3_leaf = flag
            hit.append(table[name])
        else:
            missing.append(name)
    if missing and raise_if_not_found:
        raise ValueError(f"No modules named {missing} found in model {model}")
    return hit, missing


def set_z3_leaf_modules_by_suffix(model: torch.nn.Module, module_name_suffixes: List[str], flag: bool = True,
                                  raise_if_not_found: bool = True) -> Tuple[List[torch.nn.Module], List[str]]:
    """Flag every module whose ``named_modules()`` name ends with one of the suffixes."""
    hit, missing, seen = [], [], set()
    named = list(model.named_modules())
    for suffix in module_name_suffixes:
        found = False
        for name, m in named:
            if name.endswith(suffix):
                found = True
                m._z3_leaf = flag
                if id(m) not in seen:
                    seen.add(id(m))
                    hit.append(m)
        if not found:
            missing.append(suffix)
    if missing and raise_if_not_found:
        raise ValueError(f"No modules matching suffixes {missing} found in model {model}")
    return hit, missing


def apply_zero_leaf_module_config(model: torch.nn.Module, leaf_cfg: Optional[object]) -> List[torch.nn.Module]:
    """Apply the ``zero_optimization.leaf_module`` section (classes / names / name_suffixes). Unknown entries only warn
    when the user changed that list from its default."""
    if leaf_cfg is None:
        return []
    from ..config.sections import DeepSpeedZeroLeafModuleConfig
    defaults = DeepSpeedZeroLeafModuleConfig()
    out, seen = [], set()

    def add(mods):
        for m in mods:
            if id(m) not in seen:
                seen.add(id(m))
                out.append(m)

    if leaf_cfg.classes:
        add(set_z3_leaf_modules(model, leaf_cfg.classes, raise_if_not_found=False))
    if leaf_cfg.names:
        mods, missing = set_z3_leaf_modules_by_name(model, leaf_cfg.names, raise_if_not_found=False)
        add(mods)
        if missing and leaf_cfg.names != defaults.names:
            logger.warning(f"ZeRO leaf module configuration contains unknown module names: {missing}")
    if leaf_cfg.name_suffixes:
        mods, missing = set_z3_leaf_modules_by_suffix(model, leaf_cfg.name_suffixes, raise_if_not_found=False)
        add(mods)
        if missing and leaf_cfg.name_suffixes != defaults.name_suffixes:
            logger.warning(f"ZeRO leaf module configuration contains unmatched module suffixes: {missing}")
    customized = (leaf_cfg.classes != defaults.classes or leaf_cfg.names != defaults.names
                  or leaf_cfg.name_suffixes != defaults.name_suffixes)
    if not out and customized:
        logger.warning("ZeRO leaf module configuration did not match any modules; hooks will be applied as usual")
    return out
