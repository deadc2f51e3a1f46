"""(parity: inference/v2/model_implementations/mixtral/container.py)  w1 = gate, w3 = up, w2 = down of every expert."""
from ..common_parameters import (AttentionOutputParameter, MoEGatingWeightParameter, NormParameter, UnfusedMoEGatedMLPParameter,
                                 UnfusedMoEMLP2Parameter, UnfusedQKVParameter)
from ..layer_container_base import LayerContainer
from ..llama_v2.container import Llama2NonTransformerContainer


class MixtralTransformerContainer(LayerContainer):
    qkv_w: UnfusedQKVParameter
    o_w: AttentionOutputParameter
    router_w: MoEGatingWeightParameter
    experts_gu: UnfusedMoEGatedMLPParameter
    experts_dn: UnfusedMoEMLP2Parameter
    norm1_w: NormParameter
    norm2_w: NormParameter

    PARAM_MAPPING = {
        "self_attn.q_proj.weight": "qkv_w.q_params",
        "self_attn.k_proj.weight": "qkv_w.k_params",
        "self_attn.v_proj.weight": "qkv_w.v_params",
        "self_attn.o_proj.weight": "o_w.params",
        "block_sparse_moe.gate.weight": "router_w.params",
        "block_sparse_moe.experts.*.w1.weight": "experts_gu.gating_experts",
        "block_sparse_moe.experts.*.w3.weight": "experts_gu.up_experts",
        "block_sparse_moe.experts.*.w2.weight": "experts_dn.experts",
        "input_layernorm.weight": "norm1_w.params",
        "post_attention_layernorm.weight": "norm2_w.params",
    }


class MixtralNonTransformerContainer(Llama2NonTransformerContainer):
    pass
