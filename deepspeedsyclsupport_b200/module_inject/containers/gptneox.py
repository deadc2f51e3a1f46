"""GPT-NeoX layer (parity: deepspeed/module_inject/containers/gptneox.py): fused QKV stored per head as (q, k, v) triples,
partial rotate-half rotary, parallel residual where the MLP reads its own LayerNorm of the block input."""
from ...model_implementations.transformers.ds_gpt import DeepSpeedGPTInference
from ..policy import TransformerPolicy, maybe_copy
from .base import BaseTransformerContainer
from .features.hybrid_megatron import HybridMegatronContainer
from .features.meta_tensor import MetaTensorContainer


class DS_GPTNEOXContainer(MetaTensorContainer, HybridMegatronContainer, BaseTransformerContainer):

    def __init__(self, **kwargs):
        super().__init__(**kwargs)
        self.rotate_half = True
        self.mlp_after_attn = not getattr(self.model_config, "use_parallel_residual", True)
        self.rope_theta = getattr(self.model_config, "rotary_emb_base", 10000)

    def create_module(self, config=None):
        cfg = config if config is not None else self.ds_model_config
        self.module = DeepSpeedGPTInference(cfg, mp_group=self.mp_group)
        self.module.config.scale_attention = self.scale_attention
        self.module.parallel_mlp_norm = True  # the MLP branch normalises the block input with attn_nw / attn_nb
        return self.module

    def load_params(self, module, sd, weight_quantizer, mp_replace, prefix):
        heads = self.num_attention_heads
        for dst, n in (("attn_qkvw", "attention.query_key_value.weight"), ("attn_qkvb", "attention.query_key_value.bias")):
            maybe_copy(module.attention, sd, weight_quantizer, mp_replace, dst, prefix + n, qkv=True, megatron_v2=True, split_qkv=True, heads=heads)
        for dst, n in (("attn_ow", "attention.dense.weight"), ("attn_ob", "attention.dense.bias")):
            maybe_copy(module.attention, sd, weight_quantizer, mp_replace, dst, prefix + n)
        for dst, n in (("inter_w", "mlp.dense_h_to_4h.weight"), ("inter_b", "mlp.dense_h_to_4h.bias"), ("output_w", "mlp.dense_4h_to_h.weight"),
                       ("output_b", "mlp.dense_4h_to_h.bias"), ("attn_nw", "post_attention_layernorm.weight"),
                       ("attn_nb", "post_attention_layernorm.bias")):
            maybe_copy(module.mlp, sd, weight_quantizer, mp_replace, dst, prefix + n)
        for dst, n in (("norm_w", "input_layernorm.weight"), ("norm_b", "input_layernorm.bias")):
            maybe_copy(module, sd, weight_quantizer, mp_replace, dst, prefix + n)


class GPTNEOXLayerPolicy(TransformerPolicy):
    _orig_layer_class = None
    version = 0

    def __init__(self, client_module, inference=True, megatron_v2=True, split_qkv=False):
        super().__init__(inference, megatron_v2=megatron_v2, split_qkv=split_qkv)
        self.client_module = client_module
        if GPTNEOXLayerPolicy._orig_layer_class is None:
            try:
                import transformers
                GPTNEOXLayerPolicy._orig_layer_class = transformers.models.gpt_neox.modeling_gpt_neox.GPTNeoXLayer
            except Exception:
                GPTNEOXLayerPolicy._orig_layer_class = None

    def get_hidden_heads(self):
        a = self.client_module.attention
        heads = getattr(a, "num_attention_heads", None) or a.config.num_attention_heads
        return a.query_key_value.weight.shape[1], heads, self.client_module.input_layernorm.eps, self.client_module.mlp.dense_h_to_4h.weight.shape[0]

    def attention(self, enable_training=False):
        a = self.client_module.attention
        return a.query_key_value.weight, a.query_key_value.bias, a.dense.weight, a.dense.bias

    def mlp(self, enable_training=False):
        m = self.client_module.mlp
        return m.dense_h_to_4h.weight, m.dense_h_to_4h.bias, m.dense_4h_to_h.weight, m.dense_4h_to_h.bias

    def layernorm(self):
        c = self.client_module
        return c.post_attention_layernorm.weight, c.post_attention_layernorm.bias, c.input_layernorm.weight, c.input_layernorm.bias
