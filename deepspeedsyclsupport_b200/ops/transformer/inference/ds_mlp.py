"""MLP half of the inference block (reference: ops/transformer/inference/ds_mlp.py DeepSpeedMLP over `mlp_gemm` /
`fused_gemm_gelu`): post-attention norm (attn_nw / attn_nb), up projection with fused bias + activation (gated or plain),
down projection, tensor-parallel all-reduce.  Weights are stored [out, in]."""
import torch
from torch import nn

from ....accelerator import get_accelerator
from ....utils.types import GATED_ACTIVATION_TYPES, ActivationFuncType, NormType
from ... import fused_ops as FO
from .ds_attention import _linear

_ACT = {ActivationFuncType.GELU: "gelu", ActivationFuncType.ReLU: "relu", ActivationFuncType.GATED_GELU: "gelu",
        ActivationFuncType.GATED_SILU: "silu"}


def block_norm(x, w, b, cfg):
    if cfg.norm_type == NormType.RMSNorm:
        return FO.rms_norm(x, w, cfg.epsilon)
    return FO.layer_norm(x, w, b, cfg.epsilon)


class DeepSpeedMLP(nn.Module):

    def __init__(self, config, mp_group=None, q_scales=None, q_groups=1, merge_count=1, mlp_extra_grouping=False):
        super().__init__()
        self.config = config
        self.mp_group = mp_group
        dt = torch.half if config.dtype == torch.int8 else config.dtype
        dev = get_accelerator().current_device_name()
        h = config.hidden_size
        self.gated = config.mlp_act_func_type in GATED_ACTIVATION_TYPES
        inter = config.intermediate_size // config.mp_size
        mk = lambda *s: nn.Parameter(torch.empty(*s, dtype=dt, device=dev), requires_grad=False)
        if config.set_empty_params:
            self.attn_nw = self.attn_nb = self.inter_w = self.inter_b = self.output_w = self.output_b = None
        else:
            self.attn_nw, self.attn_nb = mk(h), mk(h)
            self.inter_w, self.inter_b = mk(inter * (2 if self.gated else 1), h), mk(inter * (2 if self.gated else 1))
            self.output_w, self.output_b = mk(h, inter), mk(h)

    def mlp_only(self, hn):
        """down(act(up(hn))) with biases and the tensor-parallel all-reduce; no residual, no norm."""
        cfg = self.config
        u = _linear(hn, self.inter_w, None)
        act = _ACT[cfg.mlp_act_func_type]
        if self.gated:
            if self.inter_b is not None:
                u = u + self.inter_b
            a = FO.gated_act(u, act)
        else:
            a = FO.bias_act(u, self.inter_b, act)
        out = _linear(a, self.output_w, None)
        if self.mp_group is not None and cfg.mp_size > 1:
            from .... import comm as dist
            dist.all_reduce(out, group=self.mp_group)
        return out if self.output_b is None else out + self.output_b

    def forward(self, input, residual, residual_norm=None, bias=None):
        """input: attention output, residual: block input.
        pre-LN : x = residual + input ; returns x + mlp(norm_attn(x))
        post-LN: h = norm_attn(residual + input) ; returns h + mlp(h)   (the caller applies the final norm)"""
        cfg = self.config
        x = residual + input
        if cfg.pre_layer_norm:
            return x + self.mlp_only(block_norm(x, self.attn_nw, self.attn_nb, cfg))
        h = block_norm(x, self.attn_nw, self.attn_nb, cfg)
        return h + self.mlp_only(h)
