"""Entry points of the compression library:
   init_compression(model, ds_config[, teacher_model, mpu])  -> swap layers + switch techniques on (training time)
   redundancy_clean(model, ds_config[, mpu])                  -> bake masks / quantisation into the weights, shrink layers
   student_initialization(student, teacher, ds_config)        -> layer-reduction: copy selected teacher layers
API parity: deepspeed/compression/compress.py."""
import json
import os
import re

from . import constants as C
from .config import get_compression_config
from .helper import compression_preparation, fix_compression, is_module_compressible, recursive_getattr


def check_deepspeed_config(config):
    if isinstance(config, dict):
        return config
    if isinstance(config, str) and os.path.exists(config):
        with open(config, "r") as f:
            return json.load(f)
    raise ValueError(f"Expected a string path to an existing deepspeed config, or a dictionary. Received: {config}")


def get_module_name(group_name, model, key_word, exist_module_name, mpu=None, verbose=True):
    """Compressible modules whose qualified name matches the regex `key_word`; a module may belong to one group only."""
    found = []
    for name, module in model.named_modules():
        if re.search(key_word, name) is None or not is_module_compressible(module, mpu):
            continue
        if name in exist_module_name:
            if verbose:
                raise ValueError(f"{name} is already added to compression, please check your config file for {group_name}.")
            continue
        exist_module_name.add(name)
        found.append(name)
    return found, exist_module_name


def get_compress_methods(model, compress_methods, mpu=None):
    """-> list of [module name lists, related module name lists, {technique: params}] (one entry per config group)."""
    plan = []
    for technique, content in compress_methods.items():
        if C.LAYER_REDUCTION in technique:
            continue
        claimed = set()
        shared = content[C.SHARED_PARAMETERS]
        for group_name, group in content[C.DIFFERENT_GROUPS].items():
            names, related = [], []
            scopes = group[C.DIFFERENT_GROUPS_MODULE_SCOPE]
            related_scopes = group.get(C.DIFFERENT_GROUPS_RELATED_MODULE_SCOPE)
            for i, key_word in enumerate(scopes):
                matched, claimed = get_module_name(group_name, model, key_word, claimed, mpu=mpu)
                names.append(matched)
                if related_scopes:
                    # modules that consume the pruned module's output (e.g. Q/K/V for a pruned attention output matrix):
                    # knowing them lets redundancy_clean shrink both sides instead of only zeroing
                    related.append([get_module_name(group_name, model, rkw, set(), mpu=mpu)[0] for rkw in related_scopes[i]])
            if names:
                params = {**group[C.DIFFERENT_GROUPS_PARAMETERS], **shared}
                plan.append([names, related, {technique: params}])
    return plan


def init_compression(model, deepspeed_config, teacher_model=None, mpu=None):
    """Compress a model: replace linear/conv2d/... layers with compression-aware modules as the config describes."""
    compress_methods = get_compression_config(check_deepspeed_config(deepspeed_config))
    c_model = model.module if hasattr(model, "module") else model
    if compress_methods[C.LAYER_REDUCTION][C.LAYER_REDUCTION_ENABLED]:
        assert teacher_model is not None, "Teacher model is required for layer reduction"
        student_initialization(c_model, teacher_model, deepspeed_config)
    compression_preparation(c_model, get_compress_methods(c_model, compress_methods, mpu=mpu), mpu)
    shared = compress_methods[C.SPARSE_PRUNING][C.SHARED_PARAMETERS]
    if shared[C.SPARSE_PRUNING_ENABLED] and shared[C.SPARSE_PRUNING_METHOD] == C.SPARSE_PRUNING_METHOD_SNIP_MOMENTUM:
        from .helper import generate_pruners, register_on_step_begin
        try:
            from neural_compressor import WeightPruningConfig
        except ImportError as e:
            raise AssertionError("please ensure the neural_compressor python package is installed by pip or conda if user wants to use "
                                 "snip_momentum sparse pruning") from e
        config = WeightPruningConfig(target_sparsity=1 - shared[C.SPARSE_PRUNING_DENSE_RATIO], pattern=shared[C.SPARSE_PRUNING_BLOCK_PATTERN],
                                     pruning_frequency=shared[C.SPARSE_PRUNING_SCHEDULE_OFFSET_STRIDE],
                                     start_step=shared[C.SPARSE_PRUNING_SCHEDULE_OFFSET], end_step=shared[C.SPARSE_PRUNING_SCHEDULE_OFFSET_END],
                                     excluded_op_names=shared[C.SPARSE_PRUNING_EXCLUDED_MODULES])
        c_model.pruners = generate_pruners(config, c_model)
        register_on_step_begin(c_model)
    return model


_FIX_ORDER = [C.WEIGHT_QUANTIZATION, C.SPARSE_PRUNING, C.ROW_PRUNING, C.HEAD_PRUNING, C.CHANNEL_PRUNING, C.ACTIVATION_QUANTIZATION]


def redundancy_clean(model, deepspeed_config, mpu=None):
    """Remove the redundancy of a model: apply every enabled technique permanently (quantise first, then prune)."""
    compress_methods = get_compression_config(check_deepspeed_config(deepspeed_config))
    c_model = model.module if hasattr(model, "module") else model
    plan = sorted(get_compress_methods(c_model, compress_methods, mpu=mpu), key=lambda item: _FIX_ORDER.index(next(iter(item[2]))))
    for name_lists, related_lists, technique in plan:
        shrink = bool(related_lists)
        for i, names in enumerate(name_lists):
            masks = [fix_compression(c_model, n, technique, dim_reduction=shrink) for n in names]
            if not shrink:
                continue
            for related_names in related_lists[i]:
                for j, n in enumerate(related_names):
                    fix_compression(c_model, n, technique, mask=masks[j], dim_reduction=True)
    return model


def student_initialization(student_model, teacher_model, deepspeed_config):
    """Layer reduction: student layer i <- teacher layer teacher_layer[i] (under module_name_prefix), plus whole modules
    listed in other_module_name (embeddings, pooler, head ...)."""
    cfg = get_compression_config(check_deepspeed_config(deepspeed_config))[C.LAYER_REDUCTION]
    prefix, teacher_layer, others = cfg[C.MODULE_NAME_PREFIX], cfg[C.TEACHER_LAYER], cfg[C.OTHER_MODULE_NAME]
    pairs = [(f"{prefix}.{s}", f"{prefix}.{t}") for s, t in enumerate(teacher_layer)] + [(n, n) for n in others]
    for s_name, t_name in pairs:
        s_module, t_module = recursive_getattr(student_model, s_name), recursive_getattr(teacher_model, t_name)
        for s_param, t_param in zip(s_module.parameters(), t_module.parameters()):
            s_param.data.copy_(t_param.data)
