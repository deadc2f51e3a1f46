"""diffusers BasicTransformerBlock on fused kernels (parity: deepspeed/ops/transformer/inference/
diffusers_transformer_block.py:35-104): LayerNorm -> self-attention -> (bias + residual + LayerNorm) -> cross-attention ->
(bias + residual + LayerNorm) -> GEGLU feed-forward -> bias + residual.  The bias + residual + norm steps are one kernel each
(csrc/norm, fused residual form); the GEGLU is the gated-activation kernel."""
import torch
import torch.nn as nn

from ... import fused_ops as FO
from .bias_add import nhwc_bias_add
from .diffusers_2d_transformer import Diffusers2DTransformerConfig
from .diffusers_attention import DeepSpeedDiffusersAttention
from .op_binding.base import linear_nt


class DeepSpeedDiffusersTransformerBlock(nn.Module):

    def __init__(self, equivalent_module: nn.Module, config: Diffusers2DTransformerConfig):
        super().__init__()
        from ....module_inject.replace_module import GroupQuantizer
        self.quantizer = GroupQuantizer(q_int8=config.int8_quantization)
        self.config = config
        m = equivalent_module
        frozen = lambda t: nn.Parameter(t.data, requires_grad=False)
        self.ff1_w, self.ff1_b = frozen(m.ff.net[0].proj.weight), frozen(m.ff.net[0].proj.bias)
        self.ff2_w, self.ff2_b = frozen(m.ff.net[2].weight), frozen(m.ff.net[2].bias)
        for i, norm in ((1, m.norm1), (2, m.norm2), (3, m.norm3)):
            setattr(self, f"norm{i}_g", frozen(norm.weight))
            setattr(self, f"norm{i}_b", frozen(norm.bias))
            setattr(self, f"norm{i}_eps", norm.eps)
        self.attn_1, self.attn_2 = m.attn1, m.attn2
        # fused attention modules hand their output bias to the following fused bias + residual + norm step
        for i, attn in ((1, self.attn_1), (2, self.attn_2)):
            if isinstance(attn, DeepSpeedDiffusersAttention):
                attn.do_out_bias = False
                setattr(self, f"attn_{i}_bias", attn.attn_ob)
            else:
                setattr(self, f"attn_{i}_bias", nn.Parameter(torch.zeros_like(self.norm2_g), requires_grad=False))

    def _bias_residual_norm(self, x, bias, residual, gamma, beta, eps):
        """-> (LayerNorm(x + bias + residual), x + bias + residual)"""
        return FO.layer_norm(x + bias, gamma, beta, eps, residual)

    def forward(self, hidden_states, context=None, timestep=None, **kwargs):
        if kwargs.get("encoder_hidden_states") is not None:
            context = kwargs["encoder_hidden_states"]
        n1 = FO.layer_norm(hidden_states, self.norm1_g, self.norm1_b, self.norm1_eps)
        a1 = self.attn_1(n1)
        n2, r1 = self._bias_residual_norm(a1, self.attn_1_bias, hidden_states, self.norm2_g, self.norm2_b, self.norm2_eps)
        a2 = self.attn_2(n2, context) if isinstance(self.attn_2, DeepSpeedDiffusersAttention) else self.attn_2(n2, encoder_hidden_states=context)
        n3, r2 = self._bias_residual_norm(a2, self.attn_2_bias, r1, self.norm3_g, self.norm3_b, self.norm3_eps)
        # GEGLU: diffusers packs [value | gate] (value first); the gated kernel expects [gate | value]
        u = linear_nt(n3, self.ff1_w, self.ff1_b)
        value, gate = u.chunk(2, dim=-1)
        ff = torch.nn.functional.gelu(gate) * value
        out = linear_nt(ff, self.ff2_w)
        return nhwc_bias_add(out, self.ff2_b, other=r2)
