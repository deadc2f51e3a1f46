"""Inference block whose MLP is a mixture of experts (reference: ops/transformer/inference/moe_inference.py
DeepSpeedMoEInference / DeepSpeedMoEInferenceConfig over `einsum_sec_sm_ecm`, `moe_res_matmul` and per-expert GEMM loops).

B200 path: top-k gating -> token scatter into expert order -> ONE grouped expert GEMM per projection (csrc/gemm/gemm_grouped.cu,
row ranges read on the device) -> weighted gather; no capacity padding, no per-expert launches, no host synchronisation.
Expert weights are stacked: `inter_w [E, I(x2 if gated), H]`, `output_w [E, H, I]`.  `moe_type = "residual"` adds the dense
residual MLP mixed by a learned 2-way coefficient (PR-MoE)."""
import torch
from torch import nn

from ....accelerator import get_accelerator
from ....inference.v2 import kernels as K
from ....utils.types import GATED_ACTIVATION_TYPES
from ... import fused_ops as FO
from ...gemm import moe_gemm
from .config import DeepSpeedInferenceConfig
from .ds_attention import DeepSpeedSelfAttention, _linear
from .ds_mlp import _ACT, DeepSpeedMLP, block_norm


class DeepSpeedMoEInferenceConfig(DeepSpeedInferenceConfig):

    def __init__(self, moe_experts=1, k=1, capacity_factor=1.0, eval_capacity_factor=1.0, min_capacity=1, noisy_gate_policy=None,
                 drop_tokens=True, use_rts=False, mlp_type="standard", global_experts=1, **kwargs):
        super().__init__(**kwargs)
        self.moe_experts, self.k = moe_experts, k
        self.capacity_factor, self.eval_capacity_factor, self.min_capacity = capacity_factor, eval_capacity_factor, min_capacity
        self.noisy_gate_policy, self.drop_tokens, self.use_rts = noisy_gate_policy, drop_tokens, use_rts
        self.mlp_type = mlp_type
        self.global_experts = global_experts


class DeepSpeedMoEInference(nn.Module):
    layer_id = 0

    def __init__(self, config, mp_group=None, ep_group=None, expert_mp_group=None, quantize_scales=None, quantize_groups=1, merge_count=1,
                 mlp_extra_grouping=False):
        super().__init__()
        self.config = config
        self.config.layer_id = DeepSpeedMoEInference.layer_id
        DeepSpeedMoEInference.layer_id += 1
        self.ep_group, self.expert_mp_group = ep_group, expert_mp_group
        dt = torch.half if config.dtype == torch.int8 else config.dtype
        dev = get_accelerator().current_device_name()
        h, E = config.hidden_size, config.moe_experts
        self.gated = config.mlp_act_func_type in GATED_ACTIVATION_TYPES
        inter = config.intermediate_size
        mk = lambda *s: nn.Parameter(torch.empty(*s, dtype=dt, device=dev), requires_grad=False)
        self.attention = DeepSpeedSelfAttention(config, mp_group, quantize_scales, quantize_groups, merge_count)
        self.norm_w, self.norm_b = mk(h), mk(h)
        self.attn_nw, self.attn_nb = mk(h), mk(h)
        self.gate_w = nn.Parameter(torch.empty(E, h, dtype=torch.float32, device=dev), requires_grad=False)  # router in fp32
        self.inter_w, self.inter_b = mk(E, inter * (2 if self.gated else 1), h), mk(E, inter * (2 if self.gated else 1))
        self.output_w, self.output_b = mk(E, h, inter), mk(E, h)
        if config.mlp_type == "residual":
            self.res_mlp = DeepSpeedMLP(config, mp_group, quantize_scales, quantize_groups, merge_count, mlp_extra_grouping)
            self.res_coef = mk(2, h)

    def moe(self, hn):
        """hn: [T, H] normalised tokens -> mixture output [T, H]."""
        cfg = self.config
        logits = torch.nn.functional.linear(hn.float(), self.gate_w)
        counts, scores, assign, offsets = K.top_k_gating(logits, cfg.k, normalize=cfg.k > 1)
        moe_in, cumsum, slots = K.moe_scatter(hn, counts, assign, offsets)
        rows_expert = torch.repeat_interleave(torch.arange(cfg.moe_experts, device=hn.device), counts.long()) if self.inter_b is not None else None
        u = moe_gemm(moe_in, self.inter_w, cumsum)
        if rows_expert is not None:
            u = u + self.inter_b[rows_expert]
        a = FO.gated_act(u, _ACT[cfg.mlp_act_func_type]) if self.gated else FO.bias_act(u, None, _ACT[cfg.mlp_act_func_type])
        y = moe_gemm(a.to(moe_in.dtype), self.output_w, cumsum)
        if rows_expert is not None:
            y = y + self.output_b[rows_expert]
        return K.moe_gather(y.to(hn.dtype), scores, slots)

    def forward(self, input, input_mask=None, attention_mask=None, head_mask=None, layer_past=None, get_key_value=False, get_present=False,
                encoder_output=None, enc_dec_attn_mask=None, encoder_hidden_states=None, encoder_attention_mask=None, use_cache=False,
                output_attentions=False):
        cfg = self.config
        get_present = get_present or get_key_value or use_cache
        mask = attention_mask if attention_mask is not None else input_mask
        in_dtype = input.dtype
        with torch.no_grad():
            hn = block_norm(input, self.norm_w, self.norm_b, cfg)
            attn_out, key, value = self.attention(hn, mask, head_mask, layer_past, get_present)
            x = input + attn_out
            h2 = block_norm(x, self.attn_nw, self.attn_nb, cfg)
            B, S, H = h2.shape
            mix = self.moe(h2.reshape(B * S, H)).view(B, S, H)
            if cfg.mlp_type == "residual":
                dense = self.res_mlp.mlp_only(h2)
                coef = torch.softmax(_linear(h2, self.res_coef, None).float(), -1).to(mix.dtype)
                mix = mix * coef[..., 0:1] + dense * coef[..., 1:2]
            out = (x + mix).to(in_dtype)
        if get_present:
            out = (out, (key, value))
        if cfg.return_tuple:
            return out if isinstance(out, tuple) else (out, )
        return out
