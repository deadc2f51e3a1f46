"""Megatron-DeepSpeed GPT with MoE MLPs (parity: deepspeed/module_inject/containers/megatron_gpt_moe.py)."""
from .base_moe import BaseTransformerMoEContainer
from .features.megatron import MegatronContainer
from .megatron_gpt import MegatronLayerPolicy


class DS_MegatronGPTMoEContainer(MegatronContainer, BaseTransformerMoEContainer):
    pass


class MegatronMoELayerPolicy(MegatronLayerPolicy):
    """Experts live in `mlp.deepspeed_moe.experts.deepspeed_experts`; PR-MoE layers add a dense residual MLP + coefficient."""
    _orig_layer_class = None
    version = 0
    moe_type = "standard"
    num_experts = 1

    def __init__(self, client_module, inference=True):
        super().__init__(client_module, inference)
        if MegatronMoELayerPolicy._orig_layer_class is None:
            try:
                from megatron.model.transformer import ParallelTransformerLayer
                MegatronMoELayerPolicy._orig_layer_class = ParallelTransformerLayer
            except ImportError:
                MegatronMoELayerPolicy._orig_layer_class = None

    def get_num_experts(self):
        return self.num_experts

    def mlp(self, moe_type="standard", enable_training=False):
        moe = self.client_module.mlp.deepspeed_moe if moe_type == "standard" else self.client_module.mlp.moe.deepspeed_moe
        experts = moe.experts.deepspeed_experts
        n = len(experts)
        packed = (n, [e.dense_h_to_4h.weight for e in experts], [e.dense_h_to_4h.bias for e in experts],
                  [e.dense_4h_to_h.weight for e in experts], [e.dense_4h_to_h.bias for e in experts])
        if moe_type == "standard":
            return packed
        r = self.client_module.mlp.mlp
        return packed + (r.dense_h_to_4h.weight, r.dense_h_to_4h.bias, r.dense_4h_to_h.weight, r.dense_4h_to_h.bias,
                         self.client_module.mlp.coefficient.weight)
