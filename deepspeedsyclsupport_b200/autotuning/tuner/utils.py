"""Feature helpers for the model-based tuner (parity: tuner/utils.py)."""
import numpy as np

from ..utils import dict_to_feature, flatten, get_all_configs  # noqa: F401


def index_to_feature(p, dims):
    """index in flattened space -> mixed-radix digits"""
    feature = []
    for d in dims:
        feature.append(p % d)
        p //= d
    return feature


def feature_to_index(feature, dims):
    p = 0
    for j, k in enumerate(feature):
        p += int(np.prod(dims[:j])) * k
    return p


def dict_to_dims(tuning_space):
    dims = []
    for key, val in tuning_space.items():
        if isinstance(val, dict):
            dims.extend(dict_to_dims(val))
        elif isinstance(val, list):
            dims.append(len(val))
        else:
            dims.append(1)
    return dims


def gen_combinations(d: dict):
    import itertools
    keys, values = d.keys(), d.values()
    for v in values:
        if not isinstance(v, list):
            v = [v]
    values_choices = (gen_combinations(v) if isinstance(v, dict) else (v if isinstance(v, list) else [v]) for v in values)
    for comb in itertools.product(*values_choices):
        yield dict(zip(keys, comb))
