"""Helpers of the autotuner: nested-dict access, tuning-space expansion, experiment naming, log scanning.
(API parity with the helpers of deepspeed/autotuning/utils.py that user tooling imports.)"""
import collections.abc
import itertools
import json
import os
import re

from ..runtime.constants import GRADIENT_ACCUMULATION_STEPS, TRAIN_MICRO_BATCH_SIZE_PER_GPU
from ..utils.logging import logger

_ERR_PAT = re.compile(r"(CUDA out of memory|out of memory|OutOfMemoryError|Traceback \(most recent call last\)|RuntimeError|Error:)", re.IGNORECASE)


def search_error(filename):
    """First error-looking line of an experiment's stderr log, or None."""
    if not os.path.exists(filename):
        return "stderr.log does not exist"
    with open(filename, errors="ignore") as f:
        for line in f:
            if _ERR_PAT.search(line):
                return line.strip()
    return None


def was_interruptted(filename):
    if not os.path.exists(filename):
        return "stderr.log does not exist"
    with open(filename, errors="ignore") as f:
        return any("KeyboardInterrupt" in line for line in f)


def find_replace_str(value, replace_dict):
    if not isinstance(value, str):
        return str(value)
    for var, val in replace_dict.items():
        value = value.replace(var, str(val))
    return value


def find_replace(target, replace_dict):
    for key, value in target.items():
        if isinstance(value, str):
            target[key] = find_replace_str(value, replace_dict)
        elif isinstance(value, list):
            target[key] = [find_replace_str(v, replace_dict) for v in value]
        elif isinstance(value, dict):
            find_replace(value, replace_dict)


def get_list(val):
    return val if isinstance(val, list) else [val]


def combine_dict(d, u):
    for k, v in u.items():
        if isinstance(v, collections.abc.Mapping):
            d[k] = combine_dict(d.get(k, {}), v)
        else:
            d.setdefault(k, [])
            d[k] = list(set(get_list(d[k]) + get_list(v)))
    return d


def del_if_exists(t, d):
    """Delete key `t` anywhere in nested dict d."""
    if t in d:
        del d[t]
        return
    for v in d.values():
        if isinstance(v, collections.abc.Mapping):
            del_if_exists(t, v)


def replace_dict(d, u, ignored_keys=[]):
    """Overwrite values of d with those of u (recursively), leaving ignored_keys alone; missing keys are added."""
    if u is not None:
        for k, v in u.items():
            if k in ignored_keys:
                continue
            if v is None:
                del_if_exists(k, d)
            elif isinstance(v, collections.abc.Mapping):
                d[k] = replace_dict(d.get(k, {}), v, ignored_keys)
            else:
                d[k] = v
    return d


def get_val_by_key(d: dict, k):
    if k in d:
        return d[k]
    for v in d.values():
        if isinstance(v, dict):
            found = get_val_by_key(v, k)
            if found is not None:
                return found
    return None


def set_val_by_key(d: dict, k, vv):
    if k in d:
        d[k] = vv
    for v in d.values():
        if isinstance(v, dict):
            set_val_by_key(v, k, vv)


def fetch_hostfile(hostfile_path):
    from ..launcher.runner import fetch_hostfile as _f
    return _f(hostfile_path)


def validate_ds_config(config: dict):
    def is_false(c, key):
        return bool(c.get(key)) is False

    zero = config.get("zero_optimization", {})
    if not zero:
        return True
    stage = zero.get("stage")
    offload = False
    if stage == 1:
        return True
    if stage == 2:
        if is_false(zero, "cpu_offload") and is_false(zero, "cpu_offload_params"):
            return False
    elif stage == 3:
        offload_devices = ["cpu", "nvme"]
        if zero.get("offload_optimizer", {}).get("device") in offload_devices:
            offload = True
        if zero.get("offload_param", {}).get("device") in offload_devices:
            offload = True
    else:
        return True
    if offload:
        return False
    return True


def remove_dupe_dicts(l):
    seen, out = set(), []
    for d in l:
        key = json.dumps(d, sort_keys=True)
        if key not in seen:
            seen.add(key)
            out.append(d)
    return out


def prune_config(config, ignored_keys=[]):
    """Drop ignored keys and sub-dicts that became empty (so two configs differing only there compare equal)."""
    for k in ignored_keys:
        config.pop(k, None)
    for k in [k for k, v in config.items() if isinstance(v, dict)]:
        prune_config(config[k], ignored_keys)
        if not config[k]:
            del config[k]


def prune_configs(configs, ignored_keys=[]):
    pruned = []
    for c in configs:
        c = json.loads(json.dumps(c))
        prune_config(c, ignored_keys)
        pruned.append(c)
    return remove_dupe_dicts(pruned)


def get_tuning_keys(tuning_space: dict):
    """Keys whose value is a list with more than one candidate (searched recursively)."""
    keys = []
    for k, v in tuning_space.items():
        if isinstance(v, dict):
            keys.extend(get_tuning_keys(v))
        elif isinstance(v, list) and len(v) > 1:
            keys.append(k)
    return keys


def flatten(d, parent_key="", sep="_"):
    items = []
    for k, v in d.items():
        new_key = parent_key + sep + k if parent_key else k
        if isinstance(v, collections.abc.MutableMapping):
            items.extend(flatten(v, new_key, sep=sep).items())
        else:
            items.append((new_key, v))
    return dict(items)


def dict_to_feature(feature_dict, keys, max_value=None):
    feature = []
    for key, val in feature_dict.items():
        if key in keys and val is not None:
            if isinstance(val, bool):
                val = int(val)
            if isinstance(val, (int, float)):
                feature.append(float(val))
    return feature


def get_all_configs(tuning_space: dict, ignore_keys=None):
    """Cartesian product over every list-valued leaf -> list of concrete config dicts."""

    def expand(space):
        keys, options = [], []
        for k, v in space.items():
            if ignore_keys and k in ignore_keys:
                keys.append(k)
                options.append([v])
            elif isinstance(v, dict):
                keys.append(k)
                options.append(expand(v))
            else:
                keys.append(k)
                options.append(get_list(v))
        return [dict(zip(keys, combo)) for combo in itertools.product(*options)]

    return expand(tuning_space)


def canonical_name(config: dict, tuning_keys=None, prefix="", omit_val=False):
    """z<stage>_gas<g>_tmbspg<m>[_<key><value>...] -- stable experiment names."""
    if TRAIN_MICRO_BATCH_SIZE_PER_GPU not in (tuning_keys or []):
        tuning_keys = list(tuning_keys or []) + [TRAIN_MICRO_BATCH_SIZE_PER_GPU]
    if GRADIENT_ACCUMULATION_STEPS not in tuning_keys:
        tuning_keys = tuning_keys + [GRADIENT_ACCUMULATION_STEPS]
    tuning_keys = sorted(tuning_keys)

    def walk(cfg):
        parts = []
        for key, val in cfg.items():
            if isinstance(val, dict):
                parts.extend(walk(val))
            elif key in tuning_keys and val is not None:
                abbrev = "".join(w[0] for w in key.split("_"))
                parts.append(abbrev if omit_val else f"{abbrev}{val}")
        return parts

    name = "_".join(walk(config))
    return f"{prefix}_{name}" if prefix else name


def get_first_config(config: dict):
    """The config obtained by taking the first candidate of every list."""
    if not config:
        return None
    out = {}
    for k, v in config.items():
        if isinstance(v, dict):
            out[k] = get_first_config(v)
        elif isinstance(v, list):
            out[k] = v[0] if v else None
        else:
            out[k] = v
    return out


def write_experiments(exps: list, exps_dir: str):
    paths = []
    os.makedirs(exps_dir, exist_ok=True)
    for exp in exps:
        path = os.path.join(exps_dir, f'{exp["name"]}.json')
        with open(path, "w") as fd:
            json.dump(exp, fd)
        paths.append(path)
    return paths


def memory_to_string(n, postfix="", units=None, precision=2):
    for unit, size in (("T", 1024**4), ("G", 1024**3), ("M", 1024**2), ("K", 1024)):
        if (units is None and n >= size) or units == unit:
            return f"{round(n / size, precision)} {unit}{postfix}"
    return f"{n} {postfix}"


def number_to_string(n, postfix="", units=None, precision=2):
    for unit, size in (("B", 1e9), ("M", 1e6), ("K", 1e3)):
        if (units is None and n >= size) or units == unit:
            return f"{round(n / size, precision)} {unit}{postfix}"
    return f"{n} {postfix}"
