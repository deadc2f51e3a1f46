"""Llama decoder layer, HF format (parity: deepspeed/module_inject/containers/llama.py): RMSNorm, rotate-half rotary over the
whole head, gated SiLU MLP ([gate; up] fused), grouped-query attention through num_kv."""
import torch

from ...model_implementations.transformers.ds_llama2 import DeepSpeedLlama2Inference
from ...utils.types import ActivationFuncType, NormType
from ..policy import TransformerPolicy, maybe_copy, maybe_copy_geglu, maybe_copy_qkv
from .base import BaseTransformerContainer
from .features.gated_mlp import HybridGatedMLPContainer
from .features.meta_tensor import MetaTensorContainer
from .features.split_qkv import HybridSplitQKVContainer


class DS_LLAMAContainer(MetaTensorContainer, HybridGatedMLPContainer, HybridSplitQKVContainer, BaseTransformerContainer):

    def __init__(self, **kwargs):
        super().__init__(**kwargs)
        self.rotate_half = True
        self.mlp_after_attn = True

    def create_ds_model_config(self):
        hidden, heads, _, _ = self.policy.get_hidden_heads()
        self.rotary_dim = hidden // heads
        self.num_kv = self.policy.num_kv_heads()
        return super().create_ds_model_config()

    def create_module(self, config=None):
        cfg = config if config is not None else self.ds_model_config
        self.module = DeepSpeedLlama2Inference(cfg, mp_group=self.mp_group)
        self.module.config.scale_attention = self.scale_attention
        return self.module

    def set_lora_params(self):
        from ..policy import maybe_get_lora
        c = self.policy.client_module
        self.lora_params = [(p, maybe_get_lora(p)) for p in (c.mlp.up_proj.weight, c.mlp.gate_proj.weight, c.mlp.down_proj.weight,
                                                          c.self_attn.q_proj.weight, c.self_attn.k_proj.weight, c.self_attn.v_proj.weight,
                                                          c.self_attn.o_proj.weight)]

    def set_q_k_v(self):
        a = self.policy.client_module.self_attn
        self.qw, self.kw, self.vw = a.q_proj.weight, a.k_proj.weight, a.v_proj.weight
        self.qb, self.kb, self.vb = a.q_proj.bias, a.k_proj.bias, a.v_proj.bias

    def set_mlp_gate(self):
        m = self.policy.client_module.mlp
        self.inter_up_w, self.inter_up_b = m.up_proj.weight, None
        self.inter_gate_w, self.inter_gate_b = m.gate_proj.weight, None

    def attention_qkv_mp(self, mp_replace, reversed_dim=False):
        """GQA: q, k and v have different row counts, so the fused tensor is not three equal parts -- shard it with the
        head-aware rule instead of strided thirds."""
        from ...inference.v2.model_implementations.sharding import shard_qkv_param
        d = self.hidden_size // self.num_attention_heads
        kv = self.num_kv if self.num_kv > 0 else self.num_attention_heads
        a = self.module.attention
        for dst, src in (("attn_qkvw", self.qkvw), ("attn_qkvb", self.qkvb)):
            if src is None:
                with torch.no_grad():
                    getattr(a, dst).zero_()
                continue
            local = shard_qkv_param(src, mp_replace.gpu_index, mp_replace.mp_size, d, self.num_attention_heads, kv)
            with torch.no_grad():
                getattr(a, dst).copy_(local.to(getattr(a, dst).dtype))

    def load_params(self, module, sd, weight_quantizer, mp_replace, prefix):
        maybe_copy_qkv(module.attention, sd, weight_quantizer, mp_replace, "attn_qkvw", [prefix + f"self_attn.{p}_proj.weight" for p in "qkv"],
                       split_qkv=True)
        maybe_copy(module.attention, sd, weight_quantizer, mp_replace, "attn_ow", prefix + "self_attn.o_proj.weight")
        maybe_copy_geglu(module.mlp, sd, weight_quantizer, mp_replace, "inter_w", [prefix + "mlp.gate_proj.weight", prefix + "mlp.up_proj.weight"])
        maybe_copy(module.mlp, sd, weight_quantizer, mp_replace, "output_w", prefix + "mlp.down_proj.weight")
        maybe_copy(module.mlp, sd, weight_quantizer, mp_replace, "attn_nw", prefix + "post_attention_layernorm.weight")
        maybe_copy(module, sd, weight_quantizer, mp_replace, "norm_w", prefix + "input_layernorm.weight")


class LLAMALayerPolicy(TransformerPolicy):
    _orig_layer_class = None

    def __init__(self, client_module, inference=True):
        super().__init__(inference, mlp_act_func_type=ActivationFuncType.GATED_SILU, norm_type=NormType.RMSNorm)
        self.client_module = client_module
        try:
            import transformers
            LLAMALayerPolicy._orig_layer_class = transformers.models.llama.modeling_llama.LlamaDecoderLayer
        except Exception:
            LLAMALayerPolicy._orig_layer_class = None

    def _heads(self):
        a = self.client_module.self_attn
        cfg = getattr(a, "config", None) or TransformerPolicy.hf_model_config
        heads = getattr(a, "num_heads", None) or cfg.num_attention_heads
        kv = getattr(a, "num_key_value_heads", None) or getattr(cfg, "num_key_value_heads", heads)
        return heads, kv

    def num_kv_heads(self):
        heads, kv = self._heads()
        return kv if kv != heads else -1

    def get_hidden_heads(self):
        c = self.client_module
        eps = getattr(c.input_layernorm, "variance_epsilon", getattr(c.input_layernorm, "eps", 1e-6))
        return c.self_attn.q_proj.weight.shape[1], self._heads()[0], eps, c.mlp.gate_proj.weight.shape[0]

    def attention(self, enable_training=False):
        a = self.client_module.self_attn
        qkvw = torch.cat([a.q_proj.weight, a.k_proj.weight, a.v_proj.weight], dim=0)
        qkvb = torch.cat([a.q_proj.bias, a.k_proj.bias, a.v_proj.bias], dim=0) if a.q_proj.bias is not None else None
        return qkvw, qkvb, a.o_proj.weight, a.o_proj.bias

    def mlp(self, enable_training=False):
        m = self.client_module.mlp
        return torch.cat([m.gate_proj.weight, m.up_proj.weight], dim=0), None, m.down_proj.weight, None

    def layernorm(self):
        c = self.client_module
        return c.post_attention_layernorm.weight, None, c.input_layernorm.weight, None
