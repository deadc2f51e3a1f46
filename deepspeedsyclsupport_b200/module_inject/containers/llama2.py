"""Llama-2 in Meta's original checkpoint layout (parity: deepspeed/module_inject/containers/llama2.py): attention.wq / wk / wv /
wo, feed_forward.w1 (gate) / w3 (up) / w2 (down), attention_norm / ffn_norm."""
import torch

from ...utils.types import ActivationFuncType, NormType
from ..policy import TransformerPolicy
from .llama import DS_LLAMAContainer


class DS_LLAMA2Container(DS_LLAMAContainer):

    def set_q_k_v(self):
        a = self.policy.client_module.attention
        self.qw, self.kw, self.vw = a.wq.weight, a.wk.weight, a.wv.weight
        self.qb = self.kb = self.vb = None

    def set_mlp_gate(self):
        f = self.policy.client_module.feed_forward
        self.inter_up_w, self.inter_up_b, self.inter_gate_w, self.inter_gate_b = f.w3.weight, None, f.w1.weight, None

    def set_lora_params(self):
        self.lora_params = []


class LLAMA2LayerPolicy(TransformerPolicy):
    _orig_layer_class = None

    def __init__(self, client_module, inference=True):
        super().__init__(inference, mlp_act_func_type=ActivationFuncType.GATED_SILU, norm_type=NormType.RMSNorm)
        self.client_module = client_module
        try:
            import llama
            LLAMA2LayerPolicy._orig_layer_class = llama.model.TransformerBlock
        except Exception:
            LLAMA2LayerPolicy._orig_layer_class = None

    def num_kv_heads(self):
        a = self.client_module.attention
        kv = getattr(a, "n_kv_heads", None)
        return kv if kv and kv != self.get_hidden_heads()[1] else -1

    def get_hidden_heads(self):
        c = self.client_module
        heads = getattr(c, "n_heads", None) or c.attention.n_local_heads
        return c.attention.wq.weight.shape[1], heads, c.ffn_norm.eps, c.feed_forward.w1.weight.shape[0]

    def attention(self, enable_training=False):
        a = self.client_module.attention
        return torch.cat([a.wq.weight, a.wk.weight, a.wv.weight], dim=0), None, a.wo.weight, None

    def mlp(self, enable_training=False):
        f = self.client_module.feed_forward
        return torch.cat([f.w1.weight, f.w3.weight], dim=0), None, f.w2.weight, None

    def layernorm(self):
        c = self.client_module
        return c.ffn_norm.weight, None, c.attention_norm.weight, None
