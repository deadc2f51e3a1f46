"""BERT encoder layer (parity: deepspeed/module_inject/containers/bert.py): bidirectional, post-LN, returns a tuple."""
import torch

from ...model_implementations.transformers.ds_bert import DeepSpeedBERTInference
from ..policy import DEFAULT_INTERMEDIATE_SIZE, TransformerPolicy
from .base import BaseTransformerContainer


class DS_BERTContainer(BaseTransformerContainer):

    def __init__(self, **kwargs):
        super().__init__(**kwargs)
        self.return_tuple = True
        self.triangular_masking = False
        self.pre_layer_norm = False

    def create_module(self, config=None):
        cfg = config if config is not None else self.ds_model_config
        self.module = DeepSpeedBERTInference(cfg, mp_group=self.mp_group)
        self.module.config.scale_attention = self.scale_attention
        return self.module


class HFBertLayerPolicy(TransformerPolicy):

    def __init__(self, client_module, inference=False):
        super().__init__(inference, pre_attn_norm=False)
        self.client_module = client_module
        self.cuda_graph_supported = True
        if HFBertLayerPolicy._orig_layer_class is None:
            try:
                import transformers
                HFBertLayerPolicy._orig_layer_class = [transformers.models.bert.modeling_bert.BertLayer,
                                                       transformers.models.roberta.modeling_roberta.RobertaLayer]
            except Exception:
                HFBertLayerPolicy._orig_layer_class = None

    def get_hidden_heads(self):
        a = self.client_module.attention
        return (a.self.query.weight.shape[1], a.self.num_attention_heads, a.output.LayerNorm.eps,
                self.client_module.intermediate.dense.weight.shape[0])

    def attention(self, enable_training=False):
        s, o = self.client_module.attention.self, self.client_module.attention.output
        qkvw = torch.cat([s.query.weight, s.key.weight, s.value.weight], dim=0)
        qkvb = torch.cat([s.query.bias, s.key.bias, s.value.bias], dim=0)
        return qkvw, qkvb, o.dense.weight, o.dense.bias

    def mlp(self, enable_training=False):
        i, o = self.client_module.intermediate.dense, self.client_module.output.dense
        return i.weight, i.bias, o.weight, o.bias

    def layernorm(self):
        a, t = self.client_module.attention.output.LayerNorm, self.client_module.output.LayerNorm
        return a.weight, a.bias, t.weight, t.bias
