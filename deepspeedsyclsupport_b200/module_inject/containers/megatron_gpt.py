"""Megatron-LM GPT layer (parity: deepspeed/module_inject/containers/megatron_gpt.py).  Version 0 checkpoints name the attention
module `attention`, later ones `self_attention`; v2 stores the fused QKV per head as (q, k, v) triples."""
from ...model_implementations.transformers.ds_megatron_gpt import DeepSpeedMegatronGPTInference
from ..policy import TransformerPolicy
from .base import BaseTransformerContainer
from .features.megatron import MegatronContainer


class DS_MegatronGPTContainer(MegatronContainer, BaseTransformerContainer):

    def create_module(self, config=None):
        cfg = config if config is not None else self.ds_model_config
        self.module = DeepSpeedMegatronGPTInference(cfg, mp_group=self.mp_group)
        self.module.config.scale_attention = self.scale_attention
        if self.megatron_v2:
            self.module.config.rotate_half = True
            self.module.config.rotate_every_two = False
        return self.module


class MegatronLayerPolicy(TransformerPolicy):
    _orig_layer_class = None
    version = 0
    moe_type = "standard"
    megatron_v2 = True
    use_mup = False

    def __init__(self, client_module, inference=True):
        super().__init__(inference, megatron_v2=MegatronLayerPolicy.megatron_v2, use_mup=MegatronLayerPolicy.use_mup)
        self.client_module = client_module
        if MegatronLayerPolicy._orig_layer_class is None:
            try:
                from megatron.model.transformer import ParallelTransformerLayer
                MegatronLayerPolicy._orig_layer_class = ParallelTransformerLayer
            except ImportError:
                MegatronLayerPolicy._orig_layer_class = None

    def _attn(self):
        c = self.client_module
        return c.attention if MegatronLayerPolicy.version == 0 or not hasattr(c, "self_attention") else c.self_attention

    def get_hidden_heads(self):
        a = self._attn()
        return (a.query_key_value.weight.shape[1], a.num_attention_heads, self.client_module.input_layernorm.eps,
                self.client_module.mlp.dense_h_to_4h.weight.shape[0] if hasattr(self.client_module.mlp, "dense_h_to_4h") else -1)

    def attention(self, enable_training=False):
        a = self._attn()
        return a.query_key_value.weight, a.query_key_value.bias, a.dense.weight, a.dense.bias

    def mlp(self, moe_type="standard", enable_training=False):
        m = self.client_module.mlp
        return m.dense_h_to_4h.weight, m.dense_h_to_4h.bias, m.dense_4h_to_h.weight, m.dense_4h_to_h.bias

    def layernorm(self):
        c = self.client_module
        return c.post_attention_layernorm.weight, c.post_attention_layernorm.bias, c.input_layernorm.weight, c.input_layernorm.bias
