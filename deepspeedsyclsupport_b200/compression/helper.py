"""Model surgery for compression: swap nn.Linear / Conv2d / Embedding / BatchNorm2d (and Megatron parallel linears) for
their compression-aware versions, switch techniques on per module, and later "fix" them into the weights.
API parity: deepspeed/compression/helper.py."""
import types

import torch

from ..utils.logging import logger
from . import constants as C
from .basic_layer import (BNLayer_Compress, ColumnParallelLinear_Compress, Conv2dLayer_Compress, Embedding_Compress, LinearLayer_Compress,
                          RowParallelLinear_Compress)


def recursive_getattr(model, module_name):
    """model.a.b.0.c by dotted name."""
    out = model
    for part in module_name.split("."):
        out = getattr(out, part)
    return out


def recursive_setattr(model, module_name, module):
    parts = module_name.split(".")
    setattr(recursive_getattr(model, ".".join(parts[:-1])) if len(parts) > 1 else model, parts[-1], module)


def _adopt(new, old, extra=()):
    """Move `new` to old's device/dtype and share old's tensors."""
    new = new.to(device=old.weight.device, dtype=old.weight.dtype)
    new.weight.data = old.weight.data
    if getattr(old, "bias", None) is not None and getattr(new, "bias", None) is not None:
        new.bias.data = old.bias.data
    for name in extra:
        getattr(new, name).data = getattr(old, name).data
    return new


def _compress_version(old, mpu):
    """The compression-aware twin of `old` (or `old` itself if it already is one); None if unsupported."""
    has_bias = getattr(old, "bias", None) is not None
    if isinstance(old, (LinearLayer_Compress, Conv2dLayer_Compress, Embedding_Compress, BNLayer_Compress)):
        return old
    if isinstance(old, torch.nn.Linear):
        return _adopt(LinearLayer_Compress(old.in_features, old.out_features, bias=has_bias), old)
    if isinstance(old, torch.nn.Conv2d):
        return _adopt(Conv2dLayer_Compress(old.in_channels, old.out_channels, old.kernel_size, old.stride, old.padding, old.dilation, old.groups,
                                           has_bias, old.padding_mode), old)
    if isinstance(old, torch.nn.BatchNorm2d):
        return _adopt(BNLayer_Compress(old.num_features, old.eps, old.momentum, old.affine, old.track_running_stats), old,
                      extra=("running_mean", "running_var"))
    if isinstance(old, torch.nn.Embedding):
        return _adopt(Embedding_Compress(old.num_embeddings, old.embedding_dim, old.padding_idx, old.max_norm, old.norm_type,
                                         old.scale_grad_by_freq, old.sparse), old)
    if mpu is not None and isinstance(old, mpu.ColumnParallelLinear):
        return _adopt(ColumnParallelLinear_Compress(mpu, old.input_size, old.output_size, gather_output=old.gather_output,
                                                    skip_bias_add=old.skip_bias_add, bias=has_bias), old)
    if mpu is not None and isinstance(old, mpu.RowParallelLinear):
        return _adopt(RowParallelLinear_Compress(mpu, old.input_size, old.output_size, input_is_parallel=old.input_is_parallel,
                                                 skip_bias_add=old.skip_bias_add, bias=has_bias), old)
    return None


def _enable(module, technique, p):
    if not p[C.TECHNIQUE_ENABLED]:
        return
    if technique == C.SPARSE_PRUNING:
        module.enable_sparse_pruning(p[C.DENSE_RATIO], p[C.METHOD])
    elif technique == C.ROW_PRUNING:
        module.enable_row_pruning(p[C.DENSE_RATIO], p[C.METHOD])
    elif technique == C.HEAD_PRUNING:
        module.enable_head_pruning(p[C.DENSE_RATIO], p[C.METHOD], p[C.HEAD_PRUNING_NUM_HEADS])
    elif technique == C.CHANNEL_PRUNING:
        module.enable_channel_pruning(p[C.DENSE_RATIO], p[C.METHOD])
    elif technique == C.ACTIVATION_QUANTIZATION:
        module.enable_activation_quantization(p[C.ACTIVATION_QUANTIZE_BITS], p[C.ACTIVATION_QUANTIZE_TYPE], p[C.ACTIVATION_QUANTIZE_RANGE])
    elif technique == C.WEIGHT_QUANTIZATION:
        module.enable_weight_quantization(p[C.WEIGHT_QUANTIZE_START_BITS], p[C.WEIGHT_QUANTIZE_TARGET_BITS], p[C.WEIGHT_QUANTIZATION_PERIOD],
                                          p[C.WEIGHT_QUANTIZE_IN_FORWARD_ENABLED], p[C.WEIGHT_QUANTIZE_TYPE], p[C.WEIGHT_QUANTIZE_GROUPS])
    else:
        raise NotImplementedError("Compression technique {} is not implemented".format(technique))


def module_replacement(model, module_name, compression_technique=None, mpu=None):
    new_module = _compress_version(recursive_getattr(model, module_name), mpu)
    for technique, params in (compression_technique or {}).items():
        _enable(new_module, technique, params)
    recursive_setattr(model, module_name, new_module)


def is_module_compressible(module, mpu=None):
    ok = isinstance(module, (torch.nn.Linear, torch.nn.Conv2d, torch.nn.Embedding, torch.nn.BatchNorm2d))
    if mpu is not None:
        ok = ok or isinstance(module, (mpu.RowParallelLinear, mpu.ColumnParallelLinear))
    return ok


def compression_preparation(model, compression_technique_list, mpu):
    """First every compressible module gets its wrapper, then each (module list, related list, technique) entry switches
    its technique on."""
    for module_name in [n for n, m in model.named_modules() if is_module_compressible(m, mpu)]:
        module_replacement(model, module_name, mpu=mpu)
    for module_name_lists, _, compression_technique in compression_technique_list:
        for names in module_name_lists:
            for module_name in names:
                module_replacement(model, module_name, compression_technique)
    return model


def fix_compression(model, module_name, compression_technique, mask=None, dim_reduction=False):
    """Bake the technique into `module_name`; returns the pruning mask (for the related modules) where one exists."""
    module = recursive_getattr(model, module_name)
    for technique, p in compression_technique.items():
        on = p[C.TECHNIQUE_ENABLED]
        if technique == C.WEIGHT_QUANTIZATION and p[C.WEIGHT_QUANTIZE_IN_FORWARD_ENABLED] and on:
            return module.fix_weight_quantization()
        if technique == C.SPARSE_PRUNING and on:
            return module.fix_sparse_pruning_helper()
        if technique == C.ROW_PRUNING and (on or mask is not None):
            return module.fix_row_col_pruning_helper(mask, dim_reduction=dim_reduction)
        if technique == C.HEAD_PRUNING and (on or mask is not None):
            return module.fix_head_pruning_helper(mask, p[C.HEAD_PRUNING_NUM_HEADS], dim_reduction=dim_reduction)
        if technique == C.CHANNEL_PRUNING and (on or mask is not None):
            return module.fix_channel_pruning_helper(mask, dim_reduction=dim_reduction)


def convert_conv1d_to_linear(model, convert_type):
    """HF GPT-2 style Conv1D (weight stored [in, out]) -> nn.Linear, so the linear techniques apply."""
    c_model = model.module if hasattr(model, "module") else model
    for name in [n for n, m in c_model.named_modules() if isinstance(m, convert_type)]:
        old = recursive_getattr(c_model, name)
        new = torch.nn.Linear(old.weight.data.size(0), old.weight.data.size(1), bias=old.bias is not None)
        new.weight.data = old.weight.data.t().contiguous()
        if new.bias is not None:
            new.bias.data = old.bias.data.view(-1)
        recursive_setattr(c_model, name, new)
    return model


def generate_pruners(config, model):
    """snip_momentum sparse pruning is delegated to Intel neural_compressor (optional dependency)."""
    try:
        from neural_compressor.compression.pruner.pruners import get_pruner
        from neural_compressor.compression.pruner.utils import parse_to_prune, process_config
    except ImportError as e:
        raise AssertionError("please ensure the neural_compressor python package is installed by pip or conda if user wants to use "
                             "snip_momentum sparse pruning") from e
    assert isinstance(model, torch.nn.Module)
    pruners = []
    for info in process_config(config):
        modules = parse_to_prune(info, model)
        if modules == {}:
            logger.warning("one pruner hooks no layers, please have a check")
        pruners.append(get_pruner(info, modules))
        info["modules"] = list(modules.keys())
        info["len_of_modules"] = len(info["modules"])
        logger.info(info)
    return pruners


def register_on_step_begin(model):
    return model.register_forward_pre_hook(lambda module, inputs: [p.on_step_begin(0) for p in module.pruners] and None)


def rewrite_optimizer_step(opt: torch.optim.Optimizer):
    """Wrap opt.step with the pruners' before/after callbacks."""

    def new_step(self, closure=None):
        for pruner in getattr(self, "pruners", ()):
            pruner.on_before_optimizer_step()
        res = self.orig_step(closure) if closure is not None else self.orig_step()
        for pruner in getattr(self, "pruners", ()):
            pruner.on_after_optimizer_step()
        return res

    opt.orig_step = opt.step
    opt.step = types.MethodType(new_step, opt)
    return opt
