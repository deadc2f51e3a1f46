"""Layer policies: how to read one transformer layer of a client model (parity: deepspeed/module_inject/policy.py:14-216).

A policy wraps one instance of the client's layer class and returns its tensors in a fixed order -- attention(): (qkv weight,
qkv bias, output weight, output bias); mlp(): (up weight, up bias, down weight, down bias); layernorm(): (post-attention gamma,
beta, input gamma, beta) -- all in nn.Linear layout ([out, in]), with Q, K, V stacked [Q; K; V].  Containers
(module_inject/containers) turn that into a DeepSpeedTransformerInference block."""
from abc import ABC, abstractmethod

import torch

from ..utils.types import ActivationFuncType, NormType

transformer_param_names = ("attn_qkvw", "attn_qkvb", "attn_ow", "attn_ob", "inter_w", "inter_b", "output_w", "output_b", "attn_nw", "attn_nb",
                           "norm_w", "norm_b")


class DSPolicy(ABC):
    _orig_layer_class = None

    def __init__(self):
        self.cuda_graph_supported = False

    @abstractmethod
    def attention(self):
        raise NotImplementedError


class TransformerPolicy(DSPolicy):
    # the HF config of the model being converted; set once per model by the injection driver
    hf_model_config = None

    def __init__(self, inference=True, linear_layer=True, scale_attention=True, megatron_v2=False, use_mup=False,
                 mlp_act_func_type=ActivationFuncType.GELU, pre_attn_norm=True, use_load_prefix=False, split_qkv=True,
                 norm_type=NormType.LayerNorm):
        super().__init__()
        self.cuda_graph_supported = False
        self.inference = inference
        self.linear_layer = linear_layer  # False: weights are stored [in, out] (GPT-2's Conv1D) and get transposed
        self.scale_attention = scale_attention
        self.is_megatron_v2 = megatron_v2
        self.use_mup = use_mup
        self.mlp_act_func_type = mlp_act_func_type
        self.pre_attn_norm = pre_attn_norm
        self.use_load_prefix = use_load_prefix
        self.split_qkv = split_qkv
        self.norm_type = norm_type

    @abstractmethod
    def attention(self):
        """(qkvw, qkvb, dense_w, dense_b)"""

    @abstractmethod
    def get_hidden_heads(self):
        """(hidden size, attention heads, layer-norm epsilon, intermediate size or DEFAULT_INTERMEDIATE_SIZE)"""

    @abstractmethod
    def mlp(self):
        """(h_to_4h weight, bias, 4h_to_h weight, bias)"""

    @abstractmethod
    def layernorm(self):
        """(post-attention gamma, beta, input gamma, beta)"""


DEFAULT_INTERMEDIATE_SIZE = -1


def transpose(data):
    """[in, out] -> contiguous [out, in] (the reference does this in place with a scratch copy to bound memory; a checkpoint
    tensor here is replaced by its transposed copy and dropped)."""
    with torch.no_grad():
        return data.t().contiguous()


def _transpose(x, heads=1, mp_replace=None):
    """Megatron-v2 fused QKV is stored per head as (q, k, v) triples; regroup to [Q; K; V]."""
    heads = heads // mp_replace.mp_size if mp_replace is not None else heads
    outer = x.shape[-1] if x.dim() == 1 else x.shape[0]
    d = outer // heads // 3
    per_head = x.reshape(heads, 3, d, *x.shape[1:]) if x.dim() > 1 else x.reshape(heads, 3, d)
    parts = [per_head[:, j].reshape(heads * d, *x.shape[1:]) if x.dim() > 1 else per_head[:, j].reshape(-1) for j in range(3)]
    return torch.cat(parts, dim=0)


def maybe_copy(module, sd, weight_quantizer, mp_replace, dst_name, src_name, qkv=False, megatron_v2=False, split_qkv=False, heads=1):
    """Checkpoint-loading helper: copy sd[src_name] into getattr(module, dst_name), slicing for tensor parallelism."""
    if src_name not in sd:
        return
    tmp = sd[src_name]
    dst = getattr(module, dst_name)
    if tmp.dim() > 1 and getattr(weight_quantizer, "q_int8", False) is False and tmp.shape != dst.shape and tmp.t().shape == dst.shape:
        tmp = transpose(tmp)
    if qkv:
        if megatron_v2:
            tmp = _transpose(tmp, heads=heads).contiguous()
        dst = mp_replace.strided_copy(dst, tmp, num_splits=3) if split_qkv else mp_replace.copy(dst, tmp)
    else:
        dst = mp_replace.copy(dst, tmp)
    setattr(module, dst_name, dst)


def maybe_copy_qkv(module, sd, weight_quantizer, mp_replace, dst_name, src_names, split_qkv=False):
    """Separate q / k / v tensors -> one fused parameter."""
    if src_names[0] not in sd:
        return
    qkv = torch.cat([sd[n] for n in src_names], dim=0)
    dst = getattr(module, dst_name)
    dst = mp_replace.strided_copy(dst, qkv, num_splits=3) if split_qkv else mp_replace.copy(dst, qkv)
    setattr(module, dst_name, dst)


def maybe_copy_geglu(module, sd, weight_quantizer, mp_replace, dst_name, src_names):
    """Separate gate / up tensors -> one [gate; up] parameter, both halves cut alike."""
    if src_names[0] not in sd:
        return
    fused = torch.cat([sd[n] for n in src_names], dim=0)
    dst = mp_replace.strided_copy(getattr(module, dst_name), fused, num_splits=2)
    setattr(module, dst_name, dst)


def pack_lora_weights(p):
    return [p.lora_right_weight, p.lora_left_weight, p.lora_scaling]


def maybe_get_lora(p):
    return pack_lora_weights(p) if hasattr(p, "lora_right_weight") else []
