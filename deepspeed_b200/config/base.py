"""Typed config base: pydantic v2 model with deprecated-field migration and "auto" passthrough.

Reference behaviour: ``deepspeed/runtime/config_utils.py:17-116`` (DeepSpeedConfigModel: deprecated
fields carry ``json_schema_extra={"deprecated": True, "new_param": ...}``; values "auto" are kept as-is;
unknown keys are rejected), ``:205`` (duplicate JSON keys rejected).
"""
import collections
import json
from functools import reduce
from typing import Any

from pydantic import BaseModel, ConfigDict

from ..utils.logging import logger


class DeepSpeedConfigModel(BaseModel):
    """Base for all ds_config sections.

    * ``strict=False`` on construction drops fields whose value is the literal string ``"auto"``
      (HF integration passes those and fills them later).
    * Fields flagged ``deprecated`` are migrated onto ``new_param`` (dotted path allowed) with a warning.
    """
    model_config = ConfigDict(validate_default=True,
                              validate_assignment=True,
                              use_enum_values=True,
                              populate_by_name=True,
                              extra="forbid",
                              arbitrary_types_allowed=True,
                              protected_namespaces=())

    def __init__(self, strict=False, **data):
        if not strict:
            data = {k: v for k, v in data.items() if not (isinstance(v, str) and v == "auto" and k != "type")}
        super().__init__(**data)
        self._deprecated_fields_check()

    def _process_deprecated_field(self, dep_field):
        fields_set = self.model_fields_set
        kwargs = type(self).model_fields[dep_field].json_schema_extra or {}
        new_param_fn = kwargs.get("new_param_fn", lambda x: x)
        param_value = new_param_fn(getattr(self, dep_field))
        new_field = kwargs.get("new_param", "")
        dep_msg = kwargs.get("deprecated_msg", "")
        if dep_field in fields_set:
            logger.warning(f"Config parameter {dep_field} is deprecated" +
                           (f" use {new_field} instead" if new_field else "") + (f". {dep_msg}" if dep_msg else ""))
            if new_field and kwargs.get("set_new_param", True):
                new_leaf = new_field.split(".")[-1]
                target = self
                for part in new_field.split(".")[:-1]:
                    target = getattr(target, part)
                assert new_leaf not in target.model_fields_set or target is not self or new_leaf not in fields_set, \
                    f"Cannot provide deprecated parameter '{dep_field}' and replacing parameter '{new_field}' together"
                try:
                    setattr(target, new_leaf, param_value)
                except Exception as e:
                    logger.error(f"Tried setting value for '{new_field}' with value from deprecated '{dep_field}'")
                    raise e

    def _deprecated_fields_check(self):
        for name, field in type(self).model_fields.items():
            extra = field.json_schema_extra
            if isinstance(extra, dict) and extra.get("deprecated", False):
                self._process_deprecated_field(name)

    def get(self, key, default=None):
        return getattr(self, key, default)

    def __getitem__(self, key):
        return getattr(self, key)

    def dict(self, **kw):  # pydantic-v1 style accessor used by callers of the reference
        return self.model_dump(**kw)


def get_config_default(config, field_name):
    assert field_name in config.model_fields, f"'{field_name}' is not a field in {config}"
    assert not config.model_fields[field_name].is_required(), f"'{field_name}' is a required field"
    return config.model_fields[field_name].get_default()


class pp_int(int):
    """int that pretty-prints with thousands separators in docs/dumps (e.g. ``pp_int(5e8)``)."""

    def __new__(cls, val, custom_print_str=None):
        inst = super().__new__(cls, val)
        inst.custom_print_str = custom_print_str
        return inst

    def __repr__(self):
        return self.custom_print_str or f"{self.real:,}"


class ScientificNotationEncoder(json.JSONEncoder):
    """Large numbers as 1e+09 in config dumps."""

    def iterencode(self, o, _one_shot=False, level=0):
        indent = self.indent if self.indent is not None else 4
        prefix_close = " " * level * indent
        level += 1
        prefix = " " * level * indent
        if isinstance(o, bool):
            return "true" if o else "false"
        elif isinstance(o, (float, int)):
            return f"{o:e}" if o > 1e3 else f"{o}"
        elif isinstance(o, collections.abc.Mapping):
            x = [f'\n{prefix}"{k}": {self.iterencode(v, level=level)}' for k, v in o.items()]
            return "{" + ", ".join(x) + f"\n{prefix_close}" + "}"
        elif isinstance(o, collections.abc.Sequence) and not isinstance(o, str):
            return f"[{', '.join(map(self.iterencode, o))}]"
        return "\n, ".join(super().iterencode(o, _one_shot))


class DeepSpeedConfigObject:
    """Plain attribute bag with a JSON repr (for legacy scalar sections)."""

    def repr(self):
        return self.__dict__

    def __repr__(self):
        return json.dumps(self.__dict__, sort_keys=True, indent=4, cls=ScientificNotationEncoder, default=str)


def get_scalar_param(param_dict, param_name, param_default_value):
    return param_dict.get(param_name, param_default_value)


def get_list_param(param_dict, param_name, param_default_value):
    return param_dict.get(param_name, param_default_value)


def get_dict_param(param_dict, param_name, param_default_value):
    return param_dict.get(param_name, param_default_value)


def dict_raise_error_on_duplicate_keys(ordered_pairs):
    """``object_pairs_hook`` for json.load that rejects duplicate keys."""
    d = dict((k, v) for k, v in ordered_pairs)
    if len(d) != len(ordered_pairs):
        counter = collections.Counter([pair[0] for pair in ordered_pairs])
        keys = [key for key, value in counter.items() if value > 1]
        raise ValueError(f"Duplicate keys in DeepSpeed config: {keys}")
    return d


def get_nested(d, dotted, default=None):
    try:
        return reduce(lambda acc, k: acc[k], dotted.split("."), d)
    except (KeyError, TypeError):
        return default


def as_any(x) -> Any:
    return x
