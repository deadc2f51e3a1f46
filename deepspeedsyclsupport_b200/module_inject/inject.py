"""Swap HF BERT layers of a *training* model for the fused training layer and back (parity: deepspeed/module_inject/inject.py).
The fused layer is ops/transformer/transformer.py DeepSpeedTransformerLayer (attention on the flash kernels, fused bias-GELU /
bias-dropout-residual-norm)."""
import copy

import torch

from ..ops.transformer import DeepSpeedTransformerConfig, DeepSpeedTransformerLayer


def module_inject(layer_obj, model, config, micro_batch_size, max_seq_length, seed, preln, fp16=True):
    """Replace every `layer_obj` instance (a BertLayer-like module) in `model` by a DeepSpeedTransformerLayer holding its weights."""
    for name, child in model.named_children():
        if isinstance(child, layer_obj):
            # max_seq_length sized the reference kernel's workspace; the kernels here take the length from the input
            cuda_config = DeepSpeedTransformerConfig(batch_size=micro_batch_size, hidden_size=config.hidden_size,
                                                     intermediate_size=getattr(config, "intermediate_size", -1),
                                                     heads=config.num_attention_heads, attn_dropout_ratio=config.attention_probs_dropout_prob,
                                                     hidden_dropout_ratio=config.hidden_dropout_prob, num_hidden_layers=config.num_hidden_layers,
                                                     initializer_range=config.initializer_range, layer_norm_eps=getattr(config, "layer_norm_eps", 1e-12), seed=seed,
                                                     fp16=fp16, pre_layer_norm=preln)
            new_module = DeepSpeedTransformerLayer(cuda_config)
            a = child.attention
            qkvw = torch.cat((a.self.query.weight.data, a.self.key.weight.data, a.self.value.weight.data), 0)
            qkvb = torch.cat((a.self.query.bias.data, a.self.key.bias.data, a.self.value.bias.data), 0)
            with torch.no_grad():
                new_module.attn_qkvw.copy_(qkvw)
                new_module.attn_qkvb.copy_(qkvb)
                new_module.attn_ow.copy_(a.output.dense.weight.data)
                new_module.attn_ob.copy_(a.output.dense.bias.data)
                attn_ln = child.PostAttentionLayerNorm if preln else a.output.LayerNorm
                new_module.attn_nw.copy_(attn_ln.weight.data)
                new_module.attn_nb.copy_(attn_ln.bias.data)
                inter = child.intermediate.dense_act if preln else child.intermediate.dense
                new_module.inter_w.copy_(inter.weight.data)
                new_module.inter_b.copy_(inter.bias.data)
                new_module.output_w.copy_(child.output.dense.weight.data)
                new_module.output_b.copy_(child.output.dense.bias.data)
                out_ln = child.PreAttentionLayerNorm if preln else child.output.LayerNorm
                new_module.norm_w.copy_(out_ln.weight.data)
                new_module.norm_b.copy_(out_ln.bias.data)
            setattr(model, name, copy.deepcopy(new_module))
        else:
            module_inject(layer_obj, child, config, micro_batch_size, max_seq_length, seed, preln, fp16)
    return model
