"""CLIP text-encoder layer (parity: deepspeed/module_inject/containers/clip.py): pre-LN, causal, separate q / k / v with biases."""
import torch

from ...model_implementations.transformers.ds_gpt import DeepSpeedGPTInference
from ..policy import TransformerPolicy
from .base import BaseTransformerContainer


class DS_CLIPContainer(BaseTransformerContainer):

    def create_module(self, config=None):
        cfg = config if config is not None else self.ds_model_config
        self.module = DeepSpeedGPTInference(cfg, mp_group=self.mp_group)
        self.module.config.scale_attention = self.scale_attention
        return self.module


class HFCLIPLayerPolicy(TransformerPolicy):
    _orig_layer_class = None

    def __init__(self, client_module, inference=False):
        super().__init__(inference, pre_attn_norm=True, scale_attention=True)
        self.client_module = client_module
        self.cuda_graph_supported = True
        if HFCLIPLayerPolicy._orig_layer_class is None:
            try:
                import transformers
                HFCLIPLayerPolicy._orig_layer_class = transformers.models.clip.modeling_clip.CLIPEncoderLayer
            except Exception:
                HFCLIPLayerPolicy._orig_layer_class = None

    def get_hidden_heads(self):
        a = self.client_module.self_attn
        return a.q_proj.weight.shape[1], a.num_heads, self.client_module.layer_norm1.eps, self.client_module.mlp.fc1.weight.shape[0]

    def attention(self, enable_training=False):
        a = self.client_module.self_attn
        return (torch.cat([a.q_proj.weight, a.k_proj.weight, a.v_proj.weight], dim=0), torch.cat([a.q_proj.bias, a.k_proj.bias, a.v_proj.bias], dim=0),
                a.out_proj.weight, a.out_proj.bias)

    def mlp(self, enable_training=False):
        m = self.client_module.mlp
        return m.fc1.weight, m.fc1.bias, m.fc2.weight, m.fc2.bias

    def layernorm(self):
        c = self.client_module
        return c.layer_norm2.weight, c.layer_norm2.bias, c.layer_norm1.weight, c.layer_norm1.bias
