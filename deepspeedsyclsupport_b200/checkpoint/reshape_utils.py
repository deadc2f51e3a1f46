"""Small helpers for re-partitioning checkpoint state (parity: deepspeed/checkpoint/reshape_utils.py)."""
import os
import re
from collections import OrderedDict

import torch

from .constants import OPTIM_FILE_SUFFIX, ZERO_FILE_PREFIX


def basic_folder_validation(dir):
    assert os.path.exists(dir), f"{dir} path does not exist"
    assert os.path.isdir(dir), f"{dir} is not a folder"


def get_files_with_prefix(all_files, prefix):
    return sorted(f for f in all_files if os.path.basename(f).startswith(prefix))


def validate_files(file_list):
    for file in file_list:
        if not os.path.isfile(file):
            print(f"Error: {file} is not existent")


def get_files(dir):
    return [os.path.join(root, f) for root, _, files in os.walk(dir) for f in files]


def _natural(path):
    return [int(c) if c.isdigit() else c for c in re.split(r"(\d+)", os.path.basename(path))]


def get_zero_files(dir):
    files = get_files(dir)
    for prefix in (ZERO_FILE_PREFIX, "fp16_" + ZERO_FILE_PREFIX, "bf16_" + ZERO_FILE_PREFIX):
        hits = [f for f in get_files_with_prefix(files, prefix) if f.endswith(OPTIM_FILE_SUFFIX)]
        if hits:
            return sorted(hits, key=_natural)
    return []


def partition_data(data_list, num_partitions):
    """Split a list into `num_partitions` consecutive, equally sized chunks."""
    num_elems = len(data_list)
    assert num_elems % num_partitions == 0
    size = num_elems // num_partitions
    return [data_list[i:i + size] for i in range(0, num_elems, size)]


def _key_list_to_string(key_list):
    return ".".join(key_list)


def merge_state_dict(dict_a, dict_b, key_list):
    merged = type(dict_a)()
    for key, value in dict_b.items():
        merged[key] = merge_state(dict_a[key], value, [str(key)]) if key in dict_a else value
    return merged


def merge_state_list(list_a, list_b, key_list):
    if len(list_a) != len(list_b):
        print(f"{_key_list_to_string(key_list)}")
        raise ValueError(f"Cannot merge lists of different lengths, a = {len(list_a)} b = {len(list_b)}")
    return [merge_state(a, b, key_list) for a, b in zip(list_a, list_b)]


def merge_state(state_a, state_b, key_list=[]):
    """Recursively merge two checkpoint states: tensors are concatenated along dim 0, containers merged element-wise."""
    if type(state_a) != type(state_b):
        key_list_string = _key_list_to_string(key_list)
        print(f"key_list = {key_list_string}")
        raise ValueError(f"Cannot merge two states of types {type(state_a)} and type {type(state_b)}")
    if type(state_a) in (dict, OrderedDict):
        return merge_state_dict(state_a, state_b, key_list)
    if type(state_a) in (list, tuple):
        return type(state_a)(merge_state_list(state_a, state_b, key_list))
    if torch.is_tensor(state_a):
        return torch.cat([state_a, state_b], 0)
    return state_a
