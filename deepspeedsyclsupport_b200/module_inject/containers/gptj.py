"""GPT-J block (parity: deepspeed/module_inject/containers/gptj.py): parallel attention + MLP off one LayerNorm, interleaved
(rotate-every-two) rotary on the first rotary_dim features, no attention biases."""
import torch

from ...model_implementations.transformers.ds_gpt import DeepSpeedGPTInference
from ..policy import TransformerPolicy, maybe_copy, maybe_copy_qkv
from .base import BaseTransformerContainer
from .features.meta_tensor import MetaTensorContainer
from .features.split_qkv import HybridSplitQKVContainer


class DS_GPTJContainer(MetaTensorContainer, HybridSplitQKVContainer, BaseTransformerContainer):

    def __init__(self, **kwargs):
        super().__init__(**kwargs)
        self.rotate_every_two = True
        self.mlp_after_attn = False

    def create_module(self, config=None):
        cfg = config if config is not None else self.ds_model_config
        self.module = DeepSpeedGPTInference(cfg, mp_group=self.mp_group)
        self.module.config.scale_attention = self.scale_attention
        return self.module

    def set_lora_params(self):
        from ..policy import maybe_get_lora
        c = self.policy.client_module
        self.lora_params = [(p, maybe_get_lora(p)) for p in (c.mlp.fc_in.weight, c.mlp.fc_out.weight, c.attn.q_proj.weight, c.attn.k_proj.weight,
                                                          c.attn.v_proj.weight, c.attn.out_proj.weight)]

    def set_q_k_v(self):
        a = self.policy.client_module.attn
        self.qw, self.kw, self.vw = a.q_proj.weight, a.k_proj.weight, a.v_proj.weight
        self.qb = self.kb = self.vb = None

    def load_params(self, module, sd, weight_quantizer, mp_replace, prefix):
        maybe_copy_qkv(module.attention, sd, weight_quantizer, mp_replace, "attn_qkvw",
                       [prefix + n for n in ("attn.q_proj.weight", "attn.k_proj.weight", "attn.v_proj.weight")], split_qkv=True)
        maybe_copy(module.attention, sd, weight_quantizer, mp_replace, "attn_ow", prefix + "attn.out_proj.weight")
        for dst, n in (("inter_w", "mlp.fc_in.weight"), ("inter_b", "mlp.fc_in.bias"), ("output_w", "mlp.fc_out.weight"),
                       ("output_b", "mlp.fc_out.bias")):
            maybe_copy(module.mlp, sd, weight_quantizer, mp_replace, dst, prefix + n)
        for dst, n in (("norm_w", "ln_1.weight"), ("norm_b", "ln_1.bias")):
            maybe_copy(module, sd, weight_quantizer, mp_replace, dst, prefix + n)


class HFGPTJLayerPolicy(TransformerPolicy):
    _orig_layer_class = None

    def __init__(self, client_module, inference=True):
        super().__init__(inference, scale_attention=True)
        self.client_module = client_module
        try:
            import transformers
            HFGPTJLayerPolicy._orig_layer_class = transformers.models.gptj.modeling_gptj.GPTJBlock
        except Exception:
            HFGPTJLayerPolicy._orig_layer_class = None

    def get_hidden_heads(self):
        a = self.client_module.attn
        return a.embed_dim, a.num_attention_heads, self.client_module.ln_1.eps, self.client_module.mlp.fc_in.weight.shape[0]

    def attention(self, enable_training=False):
        a = self.client_module.attn
        return torch.cat([a.q_proj.weight, a.k_proj.weight, a.v_proj.weight], dim=0), None, a.out_proj.weight, None

    def mlp(self, enable_training=False):
        m = self.client_module.mlp
        return m.fc_in.weight, m.fc_in.bias, m.fc_out.weight, m.fc_out.bias

    def layernorm(self):
        return None, None, self.client_module.ln_1.weight, self.client_module.ln_1.bias
