"""Attention parameters (parity: inference/v2/model_implementations/common_parameters/{qkv_parameters,attn_output_parameters}.py).
Every QKV flavour ends as [all Q heads; all K heads; all V heads] before `transform_qkv_param` shards it."""
import torch

from ..parameter_base import ParameterBase


class FusedQKVParameter(ParameterBase):
    """Checkpoint already stores [Q; K; V] contiguously (GPT-2 after the Conv1D transpose)."""
    params: torch.Tensor

    def finalize(self) -> torch.Tensor:
        return self.inference_model.transform_qkv_param(self.params)


class UnfusedQKVParameter(ParameterBase):
    """Separate q / k / v projections (llama, mistral, opt, phi, gpt-j)."""
    q_params: torch.Tensor
    k_params: torch.Tensor
    v_params: torch.Tensor

    def finalize(self) -> torch.Tensor:
        return self.inference_model.transform_qkv_param(torch.cat([self.q_params, self.k_params, self.v_params], dim=0))


def megatron_qkv_reshape(param: torch.Tensor, head_size: int, n_heads: int) -> torch.Tensor:
    """[n_heads, (q k v), d, ...] interleaved per head -> [Q; K; V]."""
    assert param.shape[0] == 3 * n_heads * head_size
    per_head = param.reshape(n_heads, 3, head_size, *param.shape[1:])
    return torch.cat([per_head[:, j].reshape(n_heads * head_size, *param.shape[1:]) for j in range(3)], dim=0)


class MegatronQKVParameter(ParameterBase):
    """Per-head interleaved fused QKV (gpt-neox, falcon with one KV head per query head)."""
    params: torch.Tensor

    def finalize(self) -> torch.Tensor:
        m = self.inference_model
        return m.transform_qkv_param(megatron_qkv_reshape(self.params, m.head_size, m.n_heads_q))


def gqa_megatron_qkv_reshape(param: torch.Tensor, head_size: int, n_heads_q: int, n_heads_kv: int) -> torch.Tensor:
    """[n_kv_heads, (group query heads, k, v), d, ...] -> [Q; K; V]."""
    group = n_heads_q // n_heads_kv
    assert param.shape[0] == (n_heads_q + 2 * n_heads_kv) * head_size
    g = param.reshape(n_heads_kv, group + 2, head_size, *param.shape[1:])
    tail = param.shape[1:]
    return torch.cat([g[:, :group].reshape(n_heads_q * head_size, *tail), g[:, group].reshape(n_heads_kv * head_size, *tail),
                      g[:, group + 1].reshape(n_heads_kv * head_size, *tail)], dim=0)


class GQAMegatronQKVParameter(ParameterBase):
    """Grouped interleaved fused QKV (falcon-40b style: every KV group stores its query heads, then k, then v)."""
    params: torch.Tensor

    def finalize(self) -> torch.Tensor:
        m = self.inference_model
        return m.transform_qkv_param(gqa_megatron_qkv_reshape(self.params, m.head_size, m.n_heads_q, m.n_heads_kv))


class AttentionOutputParameter(ParameterBase):
    params: torch.Tensor

    def finalize(self) -> torch.Tensor:
        return self.inference_model.transform_attn_out_param(self.params)
