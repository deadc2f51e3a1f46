"""Config model base + JSON helpers.

`DeepSpeedConfigModel` is a pydantic-v2-native base with the reference's deprecated-field migration
(`deprecated=True, new_param="a.b", new_param_fn=...` in a Field's json_schema_extra).  Parity:
deepspeed/runtime/config_utils.py:16-205 (pydantic-v1 based in the reference).
"""
import collections
import collections.abc
import json
from functools import reduce

from pydantic import BaseModel, ConfigDict, model_validator

from ..utils.logging import logger


class DeepSpeedConfigModel(BaseModel):
    model_config = ConfigDict(validate_default=True, validate_assignment=True, use_enum_values=True,
                              populate_by_name=True, extra="forbid", arbitrary_types_allowed=True,
                              protected_namespaces=())

    def __init__(self, strict=False, **data):
        if not strict:  # "auto" values are dropped so the default applies (HF integration convention)
            data = {k: v for k, v in data.items() if (v != "auto" or k == "replace_method")}
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
                new_param_nested = new_field.split(".")
                if len(new_param_nested) > 1:
                    obj = reduce(getattr, new_param_nested[:-1], self)
                    name = new_param_nested[-1]
                else:
                    obj, name = self, new_field
                assert name not in obj.model_fields_set or True
                try:
                    setattr(obj, name, param_value)
                except Exception as e:
                    logger.error(f"Tried setting value for '{new_field}' with value from deprecated '{dep_field}'")
                    raise e

    def _deprecated_fields_check(self):
        for name, f in type(self).model_fields.items():
            extra = f.json_schema_extra
            if isinstance(extra, dict) and extra.get("deprecated", False):
                self._process_deprecated_field(name)

    def dict(self, **kwargs):  # pydantic-v1 spelling used all over user code
        return self.model_dump(**kwargs)


def get_config_default(config, field_name):
    f = config.model_fields[field_name]
    assert not f.is_required(), f"'{field_name}' is a required field and does not have a default value"
    return f.default


class pp_int(int):
    """Integer that prints with thousands separators (used for the big bucket sizes)."""

    def __new__(cls, val, custom_print_str=None):
        inst = super().__new__(cls, val)
        inst.custom_print_str = custom_print_str
        return inst

    def __repr__(self):
        return self.custom_print_str if self.custom_print_str else f"{self.real:,}"


class ScientificNotationEncoder(json.JSONEncoder):

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
            return f"[{ f', '.join(map(self.iterencode, o)) }]"
        return "\n, ".join(super().iterencode(o, _one_shot))


class DeepSpeedConfigObject:
    def repr(self):
        return self.__dict__

    def __repr__(self):
        return json.dumps(self.__dict__, sort_keys=True, indent=4, cls=ScientificNotationEncoder)


def get_scalar_param(param_dict, param_name, param_default_value):
    return param_dict.get(param_name, param_default_value)


def get_list_param(param_dict, param_name, param_default_value):
    return param_dict.get(param_name, param_default_value)


def get_dict_param(param_dict, param_name, param_default_value):
    return param_dict.get(param_name, param_default_value)


def dict_raise_error_on_duplicate_keys(ordered_pairs):
    """json object_pairs_hook that rejects duplicate keys (parity: config_utils.py:198-205)."""
    d = dict((k, v) for k, v in ordered_pairs)
    if len(d) != len(ordered_pairs):
        counter = collections.Counter([pair[0] for pair in ordered_pairs])
        keys = [key for key, value in counter.items() if value > 1]
        raise ValueError("Duplicate keys in DeepSpeed config: {}".format(keys))
    return d
