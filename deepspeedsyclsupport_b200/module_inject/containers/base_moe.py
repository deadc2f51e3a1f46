"""Container of a mixture-of-experts layer -> DeepSpeedMoEInference (parity: deepspeed/module_inject/containers/base_moe.py).
Expert weights are stacked [E, out, in] so the block's grouped GEMM serves all local experts in one launch."""
import torch

from ...ops.transformer.inference.moe_inference import DeepSpeedMoEInference, DeepSpeedMoEInferenceConfig
from .base import BaseTransformerContainer


class BaseTransformerMoEContainer(BaseTransformerContainer):

    def __init__(self, **kwargs):
        super().__init__(**kwargs)
        mc = self.model_config
        self.num_experts = getattr(mc, "num_experts", 1)
        if isinstance(self.num_experts, (list, tuple)):
            self.num_experts = self.num_experts[min(self.layer_id, len(self.num_experts) - 1)]
        self.ep_world_size = max(1, getattr(self.config.moe, "ep_size", 1) if hasattr(self.config, "moe") else 1)
        self.local_ep_size = max(1, self.num_experts // self.ep_world_size)
        self.moe_type = getattr(self.config.moe, "type", "standard") if hasattr(self.config, "moe") else "standard"
        self.top_k = getattr(mc, "topk", getattr(mc, "top_k", 1))
        self.layer_norm_eps = getattr(mc, "layernorm_epsilon", 1e-5)

    def create_ds_model_config(self):
        self.set_hidden_heads(*self.policy.get_hidden_heads())
        self.ds_model_config = DeepSpeedMoEInferenceConfig(
            hidden_size=self.hidden_size, intermediate_size=self.intermediate_size, heads=self.num_attention_heads,
            layer_norm_eps=self.layernorm_epsilon, dtype=torch.half if self.dtype == torch.int8 else self.dtype, pre_layer_norm=self.pre_layer_norm,
            mp_size=self.mp_size, return_tuple=self.return_tuple, moe_experts=self.local_ep_size, global_experts=self.num_experts,
            mlp_type=self.moe_type, k=self.top_k, scale_attn_by_inverse_layer_idx=self.scale_attn_by_inverse_layer_idx)
        return self.ds_model_config

    def create_module(self, config=None):
        cfg = config if config is not None else self.ds_model_config
        self.module = DeepSpeedMoEInference(cfg, mp_group=self.mp_group)
        return self.module

    def initialize_tensors(self, enable_training=False):
        self.set_attention(*self.policy.attention())
        self.set_layernorm(*self.policy.layernorm())
        self.set_mlp(self.config.moe.type if hasattr(self.config, "moe") else "standard")

    def set_mlp(self, config_moe_type="standard"):
        if config_moe_type == "standard":
            self.num_experts, self._h4h_w, self._h4h_b, self._4hh_w, self._4hh_b = self.policy.mlp()
        else:
            (self.num_experts, self._h4h_w, self._h4h_b, self._4hh_w, self._4hh_b, self._res_h4h_w, self._res_h4h_b, self._res_4hh_w,
             self._res_4hh_b, self._res_coef) = self.policy.mlp(config_moe_type)

    def transpose(self):
        pass  # expert and attention weights of Megatron-MoE are already [out, in]

    def apply_tensor_parallelism(self, mp_replace, **kwargs):
        self.attention_qkv_mp(mp_replace)
        self.attention_o_mp(mp_replace)
        m = self.module
        with torch.no_grad():
            for dst, src in ((m.inter_w, self._h4h_w), (m.inter_b, self._h4h_b), (m.output_w, self._4hh_w), (m.output_b, self._4hh_b)):
                if dst is not None and src is not None:
                    dst.copy_(torch.stack([t.to(dst.dtype) for t in src]) if isinstance(src, (list, tuple)) else src.to(dst.dtype))

    def copy_data_to_new_module(self):
        with torch.no_grad():
            for dst, src in ((self.module.attn_nw, self.attn_nw), (self.module.attn_nb, self.attn_nb), (self.module.norm_w, self.input_nw),
                             (self.module.norm_b, self.input_nb)):
                if dst is not None and src is not None:
                    dst.copy_(src.to(dst.dtype))
