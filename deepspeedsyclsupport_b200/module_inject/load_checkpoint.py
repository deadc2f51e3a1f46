"""Fill an already converted (kernel-injected / tensor-parallel) module tree from checkpoint state dicts (parity:
deepspeed/module_inject/load_checkpoint.py:20-290).  Used when the client model was built on the meta device: the injected
blocks have real (empty) parameters of this rank's shapes, and the checkpoint is read once, tensor by tensor."""
import re

import torch
from torch import nn

from ..model_implementations.transformers.ds_transformer import DeepSpeedTransformerInference
from .layers import EmbeddingLayer, LinearAllreduce, LinearLayer, Normalize, OPTEmbedding, RMSNormalize


def load_model_with_checkpoint(r_module, sd, mp_replace, ckpt_type, ckpt_mp_size, weight_quantizer=None, rank=0, container=None):
    """sd: list of state dicts (one per checkpoint shard that this rank reads); `container`: the architecture's container, whose
    `load_params` knows the checkpoint names of one layer."""
    sds = sd if isinstance(sd, (list, tuple)) else [sd]
    merged = {}
    for s in sds:
        merged.update(s)
    error_msgs = []

    def load(module, state_dict, prefix):
        """nn.Module._load_from_state_dict for plain layers, slicing tensors that are larger than the (TP) destination."""
        for name, p in list(module.named_parameters(recurse=False)) + list(module.named_buffers(recurse=False)):
            key = prefix + name
            if key not in state_dict:
                continue
            src = state_dict[key]
            if p.is_meta:
                p = torch.empty_like(p, device="cpu")
            dst = mp_replace.copy(p, src) if tuple(p.shape) != tuple(src.shape) else src.to(p.dtype)
            new = nn.Parameter(dst.detach().clone() if not isinstance(dst, nn.Parameter) else dst.data, requires_grad=False)
            if name in module._parameters:
                module._parameters[name] = new
            else:
                module._buffers[name] = new.data

    def load_transformer_layer(module, prefix):
        if container is None:
            raise ValueError("a container is needed to load DeepSpeedTransformerInference blocks from a checkpoint")
        container.load_params(module, merged, weight_quantizer, mp_replace, prefix)

    plain = (nn.Linear, nn.Embedding, nn.LayerNorm, EmbeddingLayer, LinearLayer, LinearAllreduce, Normalize, OPTEmbedding, RMSNormalize)

    def walk(module, prefix="", level=0):
        for name, child in module.named_children():
            child_prefix = prefix + name + "."
            if isinstance(child, DeepSpeedTransformerInference):
                load_transformer_layer(child, child_prefix)
            elif isinstance(child, plain) or type(child).__name__.endswith(("RMSNorm", "LayerNorm")):
                if any(k.startswith(child_prefix) for k in merged):
                    load(child, merged, child_prefix)
            else:
                walk(child, child_prefix, level + 1)

    walk(r_module)
    # tied input / output embeddings: reuse the loaded table
    embed = next((m for m in r_module.modules() if isinstance(m, (nn.Embedding, EmbeddingLayer)) and getattr(m, "weight", None) is not None), None)
    head = getattr(r_module, "lm_head", None)
    if embed is not None and head is not None and getattr(head, "weight", None) is not None and head.weight.is_meta:
        head.weight = embed.weight
    for m in r_module.modules():  # anything still on meta would fail at the first forward with an obscure message
        for n, p in m.named_parameters(recurse=False):
            if p.is_meta:
                error_msgs.append(f"{type(m).__name__}.{n}")
    if error_msgs:
        raise RuntimeError("parameters missing from the checkpoint: " + ", ".join(sorted(set(error_msgs))[:20]))
    del sd
    return r_module


def layer_prefixes(state_dict):
    """Distinct '<...>.<index>.' prefixes of per-layer tensors in a state dict (helps callers drive `load_params`)."""
    return sorted({m.group(1) for k in state_dict for m in [re.match(r"^(.*?\.\d+\.)", k)] if m})
