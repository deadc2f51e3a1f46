"""Post-training int8 quantisation of the projection weights of matching transformer layers (parity:
deepspeed/module_inject/module_quantize.py).  Weights become int8 tensors with a `.scale` attribute (one scale per group of
rows), the form `GroupQuantizer` produces and the quantised inference checkpoints use."""
import torch

from .replace_module import GroupQuantizer


def quantize_transformer_layer(orig_layer_impl, model, megatron=False, preln=False, group_size=64):
    """Quantise q/k/v/o and MLP weights of every `orig_layer_impl` instance in `model` in place; returns the model."""
    quantizer = GroupQuantizer(q_int8=True, group_size=group_size)

    def quantize_weight(linear):
        w = linear.weight
        if w is None or w.dtype == torch.int8 or w.dim() != 2:
            return
        groups = max(1, w.shape[0] // group_size) if w.shape[0] % group_size == 0 else 1
        quantizer.num_groups = groups
        linear.weight = quantizer.quantize(w.data)

    def quantize_fn(child):
        for m in child.modules():
            if isinstance(m, torch.nn.Linear) or type(m).__name__ == "Conv1D":
                quantize_weight(m)
        return child

    def walk(module):
        for name, child in module.named_children():
            if isinstance(child, orig_layer_impl):
                setattr(module, name, quantize_fn(child))
            else:
                walk(child)
        return module

    return walk(model)


def dequantize_weight(weight):
    """int8 weight with `.scale` -> fp32 (inverse of the above, for checks and for layers without an int8 path)."""
    scale = weight.scale
    groups = scale.numel()
    return (weight.data.float().reshape(groups, -1) * scale.reshape(groups, 1).float()).reshape(weight.shape)
