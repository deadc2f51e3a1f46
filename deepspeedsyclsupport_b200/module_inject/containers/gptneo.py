"""GPT-Neo block (parity: deepspeed/module_inject/containers/gptneo.py): unscaled attention scores, alternating global / local
(windowed) attention layers taken from config.attention_layers."""
import torch

from ...model_implementations.transformers.ds_gpt import DeepSpeedGPTInference
from ..policy import TransformerPolicy, maybe_copy, maybe_copy_qkv
from .base import BaseTransformerContainer
from .features.meta_tensor import MetaTensorContainer


class DS_GPTNEOContainer(MetaTensorContainer, BaseTransformerContainer):

    def create_module(self, config=None):
        cfg = config if config is not None else self.ds_model_config
        self.module = DeepSpeedGPTInference(cfg, mp_group=self.mp_group)
        self.module.config.scale_attention = self.scale_attention
        return self.module

    def load_params(self, module, sd, weight_quantizer, mp_replace, prefix):
        names = ("attn.attention.q_proj.weight", "attn.attention.k_proj.weight", "attn.attention.v_proj.weight", "attn.attention.out_proj.weight",
                 "attn.attention.out_proj.bias", "mlp.c_fc.weight", "mlp.c_fc.bias", "mlp.c_proj.weight", "mlp.c_proj.bias", "ln_2.weight",
                 "ln_2.bias", "ln_1.weight", "ln_1.bias")
        maybe_copy_qkv(module.attention, sd, weight_quantizer, mp_replace, "attn_qkvw", [prefix + n for n in names[:3]], split_qkv=True)
        for dst, i in (("attn_ow", 3), ("attn_ob", 4)):
            maybe_copy(module.attention, sd, weight_quantizer, mp_replace, dst, prefix + names[i])
        for dst, i in (("inter_w", 5), ("inter_b", 6), ("output_w", 7), ("output_b", 8), ("attn_nw", 9), ("attn_nb", 10)):
            maybe_copy(module.mlp, sd, weight_quantizer, mp_replace, dst, prefix + names[i])
        for dst, i in (("norm_w", 11), ("norm_b", 12)):
            maybe_copy(module, sd, weight_quantizer, mp_replace, dst, prefix + names[i])


class HFGPTNEOLayerPolicy(TransformerPolicy):
    _orig_layer_class = None

    def __init__(self, client_module, inference=True):
        super().__init__(inference, scale_attention=False)
        self.client_module = client_module
        try:
            import transformers
            HFGPTNEOLayerPolicy._orig_layer_class = transformers.models.gpt_neo.modeling_gpt_neo.GPTNeoBlock
        except Exception:
            HFGPTNEOLayerPolicy._orig_layer_class = None

    def get_hidden_heads(self):
        a = self.client_module.attn.attention
        return a.embed_dim, a.num_heads, self.client_module.ln_1.eps, self.client_module.mlp.c_fc.weight.shape[0]

    def attention(self, enable_training=False):
        a = self.client_module.attn.attention
        return torch.cat([a.q_proj.weight, a.k_proj.weight, a.v_proj.weight], dim=0), None, a.out_proj.weight, a.out_proj.bias

    def mlp(self, enable_training=False):
        m = self.client_module.mlp
        return m.c_fc.weight, m.c_fc.bias, m.c_proj.weight, m.c_proj.bias

    def layernorm(self):
        c = self.client_module
        return c.ln_2.weight, c.ln_2.bias, c.ln_1.weight, c.ln_1.bias
