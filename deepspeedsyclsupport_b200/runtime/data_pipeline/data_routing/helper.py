"""Model conversion helpers for random-LTD (parity: data_routing/helper.py:12-46)."""
from collections import OrderedDict

from .basic_layer import RandomLayerTokenDrop


def _get(model, name):
    for part in name.split("."):
        model = getattr(model, part)
    return model


def _set(model, name, module):
    parts = name.split(".")
    parent = _get(model, ".".join(parts[:-1])) if len(parts) > 1 else model
    setattr(parent, parts[-1], module)


def convert_to_random_ltd(model, convert_type):
    c_model = model.module if hasattr(model, "module") else model
    names = [name for name, module in c_model.named_modules() if isinstance(module, convert_type)]
    for name in names:
        _set(c_model, name, RandomLayerTokenDrop(_get(c_model, name)))
    model.random_ltd_initialize()
    return model


def save_without_random_ltd(model):
    c_model = model.module if hasattr(model, "module") else model
    return remove_random_ltd_state_dict(c_model.state_dict())


def remove_random_ltd_state_dict(state_dict):
    new_state_dict = OrderedDict()
    for key, value in state_dict.items():
        new_state_dict["".join(key.split(".random_ltd_layer")) if ".random_ltd_layer" in key else key] = value
    return new_state_dict
