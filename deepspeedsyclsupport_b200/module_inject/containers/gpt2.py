"""GPT-2 block (parity: deepspeed/module_inject/containers/gpt2.py): Conv1D projections ([in, out]), pre-LN, causal."""
from ...model_implementations.transformers.ds_gpt import DeepSpeedGPTInference
from ..policy import TransformerPolicy
from .base import BaseTransformerContainer


class DS_GPT2Container(BaseTransformerContainer):

    def create_module(self, config=None):
        cfg = config if config is not None else self.ds_model_config
        self.module = DeepSpeedGPTInference(cfg, mp_group=self.mp_group)
        self.module.config.scale_attention = self.scale_attention
        return self.module


class HFGPT2LayerPolicy(TransformerPolicy):
    _orig_layer_class = None

    def __init__(self, client_module, inference=True):
        super().__init__(inference, linear_layer=False)  # HF's Conv1D stores weights transposed
        self.client_module = client_module
        try:
            import transformers
            HFGPT2LayerPolicy._orig_layer_class = transformers.models.gpt2.modeling_gpt2.GPT2Block
        except Exception:
            HFGPT2LayerPolicy._orig_layer_class = None

    def get_hidden_heads(self):
        a = self.client_module.attn
        return a.embed_dim, a.num_heads, self.client_module.ln_1.eps, self.client_module.mlp.c_fc.weight.shape[1]

    def attention(self, enable_training=False):
        a = self.client_module.attn
        return a.c_attn.weight, a.c_attn.bias, a.c_proj.weight, a.c_proj.bias

    def mlp(self, enable_training=False):
        m = self.client_module.mlp
        return m.c_fc.weight, m.c_fc.bias, m.c_proj.weight, m.c_proj.bias

    def layernorm(self):
        c = self.client_module
        return c.ln_2.weight, c.ln_2.bias, c.ln_1.weight, c.ln_1.bias
