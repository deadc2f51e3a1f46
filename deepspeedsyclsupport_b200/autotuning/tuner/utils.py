"""Feature helpers for the model-based tuner (parity: tuner/utils.py): a tuning space is a nested dict whose list-valued leaves
are the dimensions; a configuration is addressed either by its flat index or by its mixed-radix digit vector."""
import itertools

import numpy as np

from ..utils import dict_to_feature, flatten, get_all_configs  # noqa: F401


def index_to_feature(p, dims):
    """flat index -> digits, least significant dimension first"""
    digits = []
    for radix in dims:
        p, digit = divmod(p, radix)
        digits.append(digit)
    return digits


def feature_to_index(feature, dims):
    weights = np.concatenate(([1], np.cumprod(dims[:-1]))) if len(dims) else np.ones(0)
    return int(sum(int(w) * int(d) for w, d in zip(weights, feature)))


def dict_to_dims(tuning_space):
    """Sizes of the dimensions in key order: a list leaf contributes its length, a scalar leaf 1, a dict its own dimensions."""
    dims = []
    for val in tuning_space.values():
        dims += dict_to_dims(val) if isinstance(val, dict) else [len(val) if isinstance(val, list) else 1]
    return dims


def gen_combinations(d: dict):
    """Every configuration of a (nested) space: the cartesian product of its leaves' choices."""

    def choices(v):
        return gen_combinations(v) if isinstance(v, dict) else (v if isinstance(v, list) else [v])

    for combo in itertools.product(*(choices(v) for v in d.values())):
        yield dict(zip(d.keys(), combo))
