"""Tensor-parallel layer replacements used by AutoTP (parity: deepspeed/module_inject/layers.py:16-59 + lm-head variant).

`LinearAllreduce` = row-parallel linear: local GEMM on the input-feature shard, then all-reduce over the TP group, then
bias.  On CUDA the GEMM is the tcgen05 kernel.  With `attach_fused` (the inference engine does it for single-node bf16 TP) GEMM
and all-reduce are ONE kernel over peer memory (parallel/fused.FusedGemmAllReduce, csrc/gemm/gemm_ar.cu); with only a symmetric
buffer attached (`attach_symm`) the reduction is the one-shot peer-memory all-reduce after the GEMM; otherwise NCCL."""
import torch
from torch import nn
from torch.nn import functional as F
from torch.nn.parameter import Parameter

from .. import comm as dist
from ..accelerator import get_accelerator


def _matmul(x, w):
    """x [..., in] @ w[out, in]^T"""
    if x.is_cuda and x.dtype == torch.bfloat16 and w.dtype == torch.bfloat16 and x.shape[-1] % 8 == 0:
        from ..ops.gemm import gemm
        x2 = x.reshape(-1, x.shape[-1])
        if not x2.is_contiguous():
            x2 = x2.contiguous()
        return gemm(x2, w).view(*x.shape[:-1], w.shape[0])
    return F.linear(x, w)


class LinearAllreduce(nn.Module):

    def __init__(self, weight, bias=None, mp_group=None):
        super().__init__()
        self.weight = weight
        self.bias = bias
        self.mp_group = mp_group
        self._symm = None
        self._fused = None

    @property
    def out_features(self):
        return self.weight.shape[0]

    @property
    def in_features(self):
        return self.weight.shape[1]

    def attach_symm(self, symm, offset=0):
        self._symm, self._symm_off = symm, offset

    def attach_fused(self, op):
        """op: a parallel.fused.FusedGemmAllReduce for this layer's out_features, shared by every layer with the same width
        (one epoch counter per flag area)."""
        assert op.n == self.weight.shape[0]
        self._fused = op

    def forward(self, input):
        fused = self._fused
        if (fused is not None and input.is_cuda and input.dtype == torch.bfloat16 and self.weight.dtype == torch.bfloat16
                and input.numel() // input.shape[-1] <= fused.max_m and input.shape[-1] % 8 == 0):
            x2 = input.reshape(-1, input.shape[-1])
            if not x2.is_contiguous():
                x2 = x2.contiguous()
            output = fused(x2, self.weight).view(*input.shape[:-1], self.weight.shape[0])
        elif self._symm is not None and input.is_cuda and input.dtype == torch.bfloat16:
            from ..parallel.fused import gemm_allreduce
            x2 = input.reshape(-1, input.shape[-1]).contiguous()
            output = gemm_allreduce(x2, self.weight, self._symm, self._symm_off).view(*input.shape[:-1], self.weight.shape[0])
        else:
            output = _matmul(input, self.weight)
            if self.mp_group is not None and dist.get_world_size(self.mp_group) > 1:
                dist.inference_all_reduce(output, group=self.mp_group)
        if self.bias is not None:
            output = output + self.bias
        return output


class LmHeadLinearAllreduce(nn.Module):

    def __init__(self, weight, rank, world_size, bias=None, mp_group=None):
        super().__init__()
        self.weight = weight
        self.bias = bias
        self.mp_group = mp_group
        self.rank = rank
        self.world_size = world_size

    def forward(self, input):
        input_shard_size = input.shape[-1] // self.world_size
        input_shard = input[..., self.rank * input_shard_size:(self.rank + 1) * input_shard_size]
        output = _matmul(input_shard.contiguous(), self.weight)
        if self.mp_group is not None and dist.get_world_size(self.mp_group) > 1:
            dist.inference_all_reduce(output, group=self.mp_group)
        if self.bias is not None:
            output = output + self.bias
        return output


class LinearLayer(nn.Module):
    """Column-parallel (output-feature shard) linear: no communication."""

    def __init__(self, weight_shape=None, dtype=torch.half, weight=None, bias=None):
        super().__init__()
        if weight is not None:
            self.weight = weight
            self.bias = bias
        else:
            self.weight = Parameter(torch.empty(weight_shape, dtype=dtype, device=get_accelerator().current_device_name()))
            self.bias = Parameter(torch.empty(weight_shape[0], dtype=dtype, device=get_accelerator().current_device_name())) \
                if bias is not None else None

    @property
    def out_features(self):
        return self.weight.shape[0]

    @property
    def in_features(self):
        return self.weight.shape[1]

    def forward(self, input):
        output = _matmul(input, self.weight)
        if self.bias is not None:
            output = output + self.bias
        return output


class Normalize(nn.Module):

    def __init__(self, dim=None, dtype=torch.float, eps=1e-5, weight=None, bias=None):
        super().__init__()
        if weight is not None:
            self.weight, self.bias = weight, bias
        else:
            self.norm = nn.LayerNorm(dim, eps=eps).to(dtype).to(get_accelerator().current_device_name())
            self.weight, self.bias = self.norm.weight, self.norm.bias
        self.eps = eps

    def forward(self, input):
        return nn.functional.layer_norm(input, input.shape[-1:], self.weight, self.bias, eps=self.eps)


class EmbeddingLayer(nn.Module):

    def __init__(self, weight_shape=None, dtype=torch.half, weight=None, bias=None):
        super().__init__()
        if weight is None:
            self.weight = Parameter(torch.empty(weight_shape[0], weight_shape[1], dtype=dtype, device=get_accelerator().current_device_name()))
        else:
            self.weight = weight

    def forward(self, input):
        return F.embedding(input, self.weight)


class OPTEmbedding(EmbeddingLayer):
    """OPT's learned positional embedding (offset 2)."""

    def __init__(self, weight_shape=None, weight=None, bias=None):
        self.offset = 2
        super().__init__(weight_shape, weight=weight)

    def forward(self, attention_mask: torch.LongTensor, past_key_values_length: int = 0):
        attention_mask = attention_mask.long()
        positions = (torch.cumsum(attention_mask, dim=1).type_as(attention_mask) * attention_mask).long() - 1
        positions = positions[:, past_key_values_length:]
        return super().forward(positions + self.offset)


class RMSNormalize(nn.Module):

    def __init__(self, dim=None, dtype=torch.float, eps=1e-5, weight=None):
        super().__init__()
        self.weight = weight if weight is not None else nn.Parameter(torch.ones(dim, dtype=dtype, device=get_accelerator().current_device_name()))
        self.eps = eps

    def forward(self, hidden_states):
        from ..ops.fused_ops import rms_norm
        if hidden_states.is_cuda and hidden_states.dtype == self.weight.dtype and hidden_states.shape[-1] % 8 == 0:
            return rms_norm(hidden_states.contiguous(), self.weight, self.eps)
        variance = hidden_states.to(torch.float32).pow(2).mean(-1, keepdim=True)
        hidden_states = hidden_states * torch.rsqrt(variance + self.eps)
        return hidden_states.to(self.weight.dtype) * self.weight
