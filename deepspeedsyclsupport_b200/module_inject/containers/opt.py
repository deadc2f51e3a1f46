"""OPT decoder layer (parity: deepspeed/module_inject/containers/opt.py): separate q / k / v with biases, ReLU (or GELU) MLP,
pre- or post-LN by config.do_layer_norm_before."""
import torch

from ...model_implementations.transformers.ds_opt import DeepSpeedOPTInference
from ...utils.types import ActivationFuncType
from ..policy import TransformerPolicy, maybe_copy, maybe_copy_qkv
from .base import BaseTransformerContainer
from .features.meta_tensor import MetaTensorContainer
from .features.split_qkv import HybridSplitQKVContainer


class DS_OPTContainer(MetaTensorContainer, HybridSplitQKVContainer, BaseTransformerContainer):

    def create_module(self, config=None):
        cfg = config if config is not None else self.ds_model_config
        self.module = DeepSpeedOPTInference(cfg, mp_group=self.mp_group)
        self.module.config.scale_attention = self.scale_attention
        return self.module

    def set_lora_params(self):
        from ..policy import maybe_get_lora
        c = self.policy.client_module
        self.lora_params = [(p, maybe_get_lora(p)) for p in (c.fc1.weight, c.fc2.weight, c.self_attn.q_proj.weight, c.self_attn.k_proj.weight,
                                                          c.self_attn.v_proj.weight, c.self_attn.out_proj.weight)]

    def set_q_k_v(self):
        a = self.policy.client_module.self_attn
        self.qw, self.qb, self.kw, self.kb, self.vw, self.vb = (a.q_proj.weight, a.q_proj.bias, a.k_proj.weight, a.k_proj.bias, a.v_proj.weight,
                                                                a.v_proj.bias)

    def load_params(self, module, sd, weight_quantizer, mp_replace, prefix):
        for dst, kind in (("attn_qkvw", "weight"), ("attn_qkvb", "bias")):
            maybe_copy_qkv(module.attention, sd, weight_quantizer, mp_replace, dst,
                           [prefix + f"self_attn.{p}_proj.{kind}" for p in "qkv"], split_qkv=True)
        for dst, n in (("attn_ow", "self_attn.out_proj.weight"), ("attn_ob", "self_attn.out_proj.bias")):
            maybe_copy(module.attention, sd, weight_quantizer, mp_replace, dst, prefix + n)
        for dst, n in (("inter_w", "fc1.weight"), ("inter_b", "fc1.bias"), ("output_w", "fc2.weight"), ("output_b", "fc2.bias"),
                       ("attn_nw", "final_layer_norm.weight"), ("attn_nb", "final_layer_norm.bias")):
            maybe_copy(module.mlp, sd, weight_quantizer, mp_replace, dst, prefix + n)
        for dst, n in (("norm_w", "self_attn_layer_norm.weight"), ("norm_b", "self_attn_layer_norm.bias")):
            maybe_copy(module, sd, weight_quantizer, mp_replace, dst, prefix + n)


class HFOPTLayerPolicy(TransformerPolicy):
    _orig_layer_class = None

    def __init__(self, client_module, inference=True, use_load_prefix=True):
        super().__init__(inference, linear_layer=True, pre_attn_norm=True, use_load_prefix=use_load_prefix)
        self.client_module = client_module
        try:
            import transformers
            HFOPTLayerPolicy._orig_layer_class = transformers.models.opt.modeling_opt.OPTDecoderLayer
        except Exception:
            HFOPTLayerPolicy._orig_layer_class = None
        act = getattr(TransformerPolicy.hf_model_config, "activation_function", "relu") if TransformerPolicy.hf_model_config is not None else "relu"
        if act == "relu":
            self.mlp_act_func_type = ActivationFuncType.ReLU
        elif act in ("gelu", "gelu_new"):
            self.mlp_act_func_type = ActivationFuncType.GELU
        else:
            raise ValueError(f"Unsupported activation function for OPT: {act}")

    def get_hidden_heads(self):
        a = self.client_module.self_attn
        return a.embed_dim, a.num_heads, self.client_module.self_attn_layer_norm.eps, self.client_module.fc1.weight.shape[0]

    def attention(self, enable_training=False):
        a = self.client_module.self_attn
        return (torch.cat([a.q_proj.weight, a.k_proj.weight, a.v_proj.weight], dim=0), torch.cat([a.q_proj.bias, a.k_proj.bias, a.v_proj.bias], dim=0),
                a.out_proj.weight, a.out_proj.bias)

    def mlp(self, enable_training=False):
        c = self.client_module
        return c.fc1.weight, c.fc1.bias, c.fc2.weight, c.fc2.bias

    def layernorm(self):
        c = self.client_module
        return c.final_layer_norm.weight, c.final_layer_norm.bias, c.self_attn_layer_norm.weight, c.self_attn_layer_norm.bias
