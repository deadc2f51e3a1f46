"""Configuration of one inference transformer block (reference: ops/transformer/inference/config.py DeepSpeedInferenceConfig).
Same field names and defaults, so policies / containers written against the reference construct it unchanged."""
import json

import torch

from ....utils.types import ActivationFuncType, NormType

_DEFAULTS = dict(
    hidden_size=-1, intermediate_size=-1, heads=-1, num_hidden_layers=-1, layer_norm_eps=1e-12, local_rank=-1, mp_size=1,
    dtype=torch.float16, pre_layer_norm=True, norm_type=NormType.LayerNorm, stochastic_mode=False, scale_attention=True,
    triangular_masking=True, local_attention=False, window_size=256, rotary_dim=-1, rotate_half=False, rotate_every_two=True,
    return_tuple=True, mlp_after_attn=True, mlp_act_func_type=ActivationFuncType.GELU, training_mp_size=1, bigscience_bloom=False,
    max_out_tokens=1024, min_out_tokens=1, enable_qkv_quantization=False, use_mup=False, scale_attn_by_inverse_layer_idx=False,
    return_single_tuple=False, set_empty_params=False, transposed_mode=False, use_triton=False, triton_autotune=False, num_kv=-1,
    rope_theta=10000)


class TransformerConfig(object):

    def __init__(self, hidden_size, intermediate_size, heads, num_hidden_layers):
        self.layer_id = -1
        self.hidden_size, self.intermediate_size, self.heads, self.num_hidden_layers = hidden_size, intermediate_size, heads, num_hidden_layers


class DeepSpeedInferenceConfig(TransformerConfig):

    def __init__(self, **kwargs):
        unknown = set(kwargs) - set(_DEFAULTS)
        if unknown:
            raise TypeError(f"DeepSpeedInferenceConfig got unexpected arguments {sorted(unknown)}")
        v = dict(_DEFAULTS, **kwargs)
        inter = v["intermediate_size"] if v["intermediate_size"] > 0 else 4 * v["hidden_size"]
        super().__init__(v["hidden_size"], inter, v["heads"], v["num_hidden_layers"])
        for k, val in v.items():
            if k in ("hidden_size", "intermediate_size", "heads", "num_hidden_layers"):
                continue
            setattr(self, "epsilon" if k == "layer_norm_eps" else k, val)
        self.specialized_mode = False

    @classmethod
    def from_dict(cls, json_object):
        cfg = cls()
        cfg.__dict__.update(json_object)
        return cfg

    @classmethod
    def from_json_file(cls, json_file):
        with open(json_file, "r", encoding="utf-8") as f:
            return cls.from_dict(json.load(f))
