"""DistilBERT TransformerBlock (parity: deepspeed/module_inject/containers/distil_bert.py)."""
import torch

from ...model_implementations.transformers.ds_bert import DeepSpeedBERTInference
from ..policy import TransformerPolicy
from .base import BaseTransformerContainer


class DS_DistilBERTContainer(BaseTransformerContainer):

    def __init__(self, **kwargs):
        super().__init__(**kwargs)
        self.triangular_masking = False
        self.return_single_tuple = True
        self.pre_layer_norm = False

    def create_module(self, config=None):
        cfg = config if config is not None else self.ds_model_config
        self.module = DeepSpeedBERTInference(cfg, mp_group=self.mp_group)
        self.module.config.scale_attention = self.scale_attention
        return self.module


class HFDistilBertLayerPolicy(TransformerPolicy):
    _orig_layer_class = None

    def __init__(self, client_module, inference=False, preln=False):
        super().__init__(inference, pre_attn_norm=preln)
        self.client_module = client_module
        self.cuda_graph_supported = True
        if HFDistilBertLayerPolicy._orig_layer_class is None:
            try:
                import transformers
                HFDistilBertLayerPolicy._orig_layer_class = [transformers.models.distilbert.modeling_distilbert.TransformerBlock]
            except Exception:
                HFDistilBertLayerPolicy._orig_layer_class = None

    def get_hidden_heads(self):
        a = self.client_module.attention
        return a.q_lin.weight.shape[1], a.n_heads, self.client_module.sa_layer_norm.eps, self.client_module.ffn.lin1.weight.shape[0]

    def attention(self, enable_training=False):
        a = self.client_module.attention
        return (torch.cat([a.q_lin.weight, a.k_lin.weight, a.v_lin.weight], dim=0), torch.cat([a.q_lin.bias, a.k_lin.bias, a.v_lin.bias], dim=0),
                a.out_lin.weight, a.out_lin.bias)

    def mlp(self, enable_training=False):
        f = self.client_module.ffn
        return f.lin1.weight, f.lin1.bias, f.lin2.weight, f.lin2.bias

    def layernorm(self):
        a, t = self.client_module.sa_layer_norm, self.client_module.output_layer_norm
        return a.weight, a.bias, t.weight, t.bias
