"""BLOOM block (parity: deepspeed/module_inject/containers/bloom.py): fused QKV per head as (q, k, v) triples, ALiBi position
biases handed to the block by the model's forward (`alibi=`), no rotary."""
from ...model_implementations.transformers.ds_bloom import DeepSpeedBloomInference
from ..policy import TransformerPolicy, _transpose, maybe_copy
from .base import BaseTransformerContainer
from .features.hybrid_engine import HybridEngineContainer
from .features.meta_tensor import MetaTensorContainer


class DS_BloomContainer(MetaTensorContainer, HybridEngineContainer, BaseTransformerContainer):

    def __init__(self, **kwargs):
        super().__init__(**kwargs)
        self.bigscience_bloom = True
        self.triangular_masking = False  # the model passes its own causal + ALiBi mask

    def create_module(self, config=None):
        cfg = config if config is not None else self.ds_model_config
        self.module = DeepSpeedBloomInference(cfg, mp_group=self.mp_group)
        self.module.config.scale_attention = self.scale_attention
        return self.module

    def transpose(self):
        super().transpose()
        self.qkvw = _transpose(self.qkvw.data, heads=self.num_attention_heads)
        if self.qkvb is not None:
            self.qkvb = _transpose(self.qkvb.data, heads=self.num_attention_heads)

    def set_lora_params(self):
        from ..policy import maybe_get_lora
        c = self.policy.client_module
        self.lora_params = [(p, maybe_get_lora(p)) for p in (c.mlp.dense_h_to_4h.weight, c.mlp.dense_4h_to_h.weight,
                                                          c.self_attention.query_key_value.weight, c.self_attention.dense.weight)]

    def load_params(self, module, sd, weight_quantizer, mp_replace, prefix):
        heads = self.num_attention_heads
        for dst, n in (("attn_qkvw", "self_attention.query_key_value.weight"), ("attn_qkvb", "self_attention.query_key_value.bias")):
            maybe_copy(module.attention, sd, weight_quantizer, mp_replace, dst, prefix + n, qkv=True, megatron_v2=True, split_qkv=True, heads=heads)
        for dst, n in (("attn_ow", "self_attention.dense.weight"), ("attn_ob", "self_attention.dense.bias")):
            maybe_copy(module.attention, sd, weight_quantizer, mp_replace, dst, prefix + n)
        for dst, n in (("inter_w", "mlp.dense_h_to_4h.weight"), ("inter_b", "mlp.dense_h_to_4h.bias"), ("output_w", "mlp.dense_4h_to_h.weight"),
                       ("output_b", "mlp.dense_4h_to_h.bias"), ("attn_nw", "post_attention_layernorm.weight"),
                       ("attn_nb", "post_attention_layernorm.bias")):
            maybe_copy(module.mlp, sd, weight_quantizer, mp_replace, dst, prefix + n)
        for dst, n in (("norm_w", "input_layernorm.weight"), ("norm_b", "input_layernorm.bias")):
            maybe_copy(module, sd, weight_quantizer, mp_replace, dst, prefix + n)


class BLOOMLayerPolicy(TransformerPolicy):
    _orig_layer_class = None

    def __init__(self, client_module, inference=True, use_load_prefix=True, split_qkv=False):
        super().__init__(inference, linear_layer=True, use_load_prefix=use_load_prefix, split_qkv=split_qkv)
        self.client_module = client_module
        try:
            import transformers
            BLOOMLayerPolicy._orig_layer_class = transformers.models.bloom.modeling_bloom.BloomBlock
        except Exception:
            BLOOMLayerPolicy._orig_layer_class = None

    def get_hidden_heads(self):
        a = self.client_module.self_attention
        return a.hidden_size, a.num_heads, self.client_module.input_layernorm.eps, self.client_module.mlp.dense_h_to_4h.weight.shape[0]

    def attention(self, enable_training=False):
        a = self.client_module.self_attention
        return a.query_key_value.weight, a.query_key_value.bias, a.dense.weight, a.dense.bias

    def mlp(self, enable_training=False):
        m = self.client_module.mlp
        return m.dense_h_to_4h.weight, m.dense_h_to_4h.bias, m.dense_4h_to_h.weight, m.dense_4h_to_h.bias

    def layernorm(self):
        c = self.client_module
        return c.post_attention_layernorm.weight, c.post_attention_layernorm.bias, c.input_layernorm.weight, c.input_layernorm.bias
