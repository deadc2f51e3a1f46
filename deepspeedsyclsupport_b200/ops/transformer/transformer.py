"""DeepSpeedTransformerLayer: the fused BERT-style encoder layer (pre- or post-LayerNorm) used for BERT pre-training.

Reference: a monolithic C++/CUDA layer object per layer id (csrc/transformer/ds_transformer_cuda.cpp + 9 .cu files,
deepspeed/ops/transformer/transformer.py:296-412) with hand-managed activation buffers.  Here the layer is composed of
this framework's kernels and autograd does the bookkeeping:
    QKV / attn-out / FF1 / FF2 projections : tcgen05 GEMM (ops/gemm.linear; weight grads accumulate into ZeRO buffers)
    LayerNorm (+ residual add)             : fused norm kernel (ops/fused_ops.layer_norm)
    bias + GELU                            : fused elementwise kernel forward, recomputed in backward
                                             (= the reference's `gelu_checkpoint` behaviour, always on)
    attention                              : fused SDPA with the additive padding mask, dropout inside the kernel
Parameter names (attn_qkvw, attn_qkvb, attn_ow, ..., norm_b), config fields and the forward signature follow the
reference so checkpoints and model code carry over.  `normalize_invertible`, `attn_dropout_checkpoint` and
`stochastic_mode` are accepted; they were memory/perf switches of the CUDA object and do not change results."""
import json
import math

import torch
import torch.nn.functional as F
from torch import nn

from ...accelerator import get_accelerator
from .. import fused_ops
from ..gemm import linear


class TransformerConfig(object):

    def __init__(self, batch_size, hidden_size, intermediate_size, heads, attn_dropout_ratio, hidden_dropout_ratio, num_hidden_layers, initializer_range):
        self.layer_id = -1
        self.batch_size = batch_size
        self.hidden_size = hidden_size
        self.intermediate_size = intermediate_size
        self.heads = heads
        self.attn_dropout_ratio = attn_dropout_ratio
        self.hidden_dropout_ratio = hidden_dropout_ratio
        self.num_hidden_layers = num_hidden_layers
        self.initializer_range = initializer_range


class DeepSpeedTransformerConfig(TransformerConfig):

    def __init__(self, batch_size=-1, hidden_size=-1, intermediate_size=-1, heads=-1, attn_dropout_ratio=-1, hidden_dropout_ratio=-1,
                 num_hidden_layers=-1, initializer_range=-1, layer_norm_eps=1e-12, local_rank=-1, seed=-1, fp16=False, pre_layer_norm=True,
                 normalize_invertible=False, gelu_checkpoint=False, adjust_init_range=True, attn_dropout_checkpoint=False, stochastic_mode=False,
                 return_tuple=False, training=True, bf16=False):
        super().__init__(batch_size, hidden_size, (intermediate_size if intermediate_size > 0 else 4 * hidden_size), heads, attn_dropout_ratio,
                         hidden_dropout_ratio, num_hidden_layers, initializer_range)
        self.fp16 = fp16
        self.bf16 = bf16
        self.pre_layer_norm = pre_layer_norm
        self.local_rank = local_rank
        self.seed = seed
        self.normalize_invertible = normalize_invertible
        self.gelu_checkpoint = gelu_checkpoint
        self.adjust_init_range = adjust_init_range
        self.test_gemm = False
        self.layer_norm_eps = layer_norm_eps
        self.training = training
        self.is_grad_enabled = True
        self.attn_dropout_checkpoint = attn_dropout_checkpoint
        self.stochastic_mode = stochastic_mode
        self.return_tuple = return_tuple

    @classmethod
    def from_dict(cls, json_object):
        config = DeepSpeedTransformerConfig()
        for key, value in json_object.items():
            config.__dict__[key] = value
        return config

    @classmethod
    def from_json_file(cls, json_file):
        with open(json_file, "r", encoding="utf-16") as reader:
            return cls.from_dict(json.loads(reader.read()))


class _BiasGelu(torch.autograd.Function):
    """gelu(x + b): fused kernel forward; backward recomputes the pre-activation instead of saving it."""

    @staticmethod
    def forward(ctx, x, bias):
        ctx.save_for_backward(x, bias)
        return fused_ops.bias_act(x, bias, act="gelu")

    @staticmethod
    def backward(ctx, dy):
        x, bias = ctx.saved_tensors
        v = x.float() + bias.float()
        k = math.sqrt(2.0 / math.pi)
        inner = k * (v + 0.044715 * v**3)
        t = torch.tanh(inner)
        dv = 0.5 * (1 + t) + 0.5 * v * (1 - t * t) * k * (1 + 3 * 0.044715 * v * v)
        dx = (dy.float() * dv)
        return dx.to(x.dtype), dx.reshape(-1, dx.shape[-1]).sum(0).to(bias.dtype)


class DeepSpeedTransformerLayer(nn.Module):
    layer_id = 0

    def __init__(self, config, initial_weights=None, initial_biases=None):
        super().__init__()
        self.config = config
        self.config.layer_id = DeepSpeedTransformerLayer.layer_id
        DeepSpeedTransformerLayer.layer_id += 1
        if self.config.local_rank >= 0 and get_accelerator().is_available():
            get_accelerator().set_device(self.config.local_rank)
        H, I = config.hidden_size, config.intermediate_size
        if initial_weights is None and initial_biases is None:
            self.attn_qkvw = nn.Parameter(torch.empty(H * 3, H))
            self.attn_qkvb = nn.Parameter(torch.empty(H * 3))
            self.attn_ow = nn.Parameter(torch.empty(H, H))
            self.attn_ob = nn.Parameter(torch.empty(H))
            self.attn_nw = nn.Parameter(torch.empty(H))
            self.attn_nb = nn.Parameter(torch.empty(H))
            self.inter_w = nn.Parameter(torch.empty(I, H))
            self.inter_b = nn.Parameter(torch.empty(I))
            self.output_w = nn.Parameter(torch.empty(H, I))
            self.output_b = nn.Parameter(torch.empty(H))
            self.norm_w = nn.Parameter(torch.empty(H))
            self.norm_b = nn.Parameter(torch.empty(H))
            self.init_transformer_weights(self.config.adjust_init_range)
        else:
            # unit-test path: adopt the weights of a reference BERT layer (q, k, v, attn_out, attn_norm, inter, output, norm)
            self.attn_qkvw = nn.Parameter(torch.cat([initial_weights[i].data for i in range(3)]))
            self.attn_qkvb = nn.Parameter(torch.cat([initial_biases[i].data for i in range(3)]) if initial_biases[0] is not None else torch.zeros(H * 3))
            self.attn_ow, self.attn_ob = initial_weights[3], initial_biases[3]
            self.attn_nw, self.attn_nb = initial_weights[4], initial_biases[4]
            self.inter_w, self.inter_b = initial_weights[5], initial_biases[5]
            self.output_w, self.output_b = initial_weights[6], initial_biases[6]
            self.norm_w, self.norm_b = initial_weights[7], initial_biases[7]

    def init_transformer_weights(self, adjust_init_range=False):
        std = self.config.initializer_range
        output_std = std / math.sqrt(2.0 * self.config.num_hidden_layers) if adjust_init_range else std
        self.attn_qkvw.data.normal_(mean=0.0, std=std)
        self.attn_ow.data.normal_(mean=0.0, std=output_std)
        self.inter_w.data.normal_(mean=0.0, std=std)
        self.output_w.data.normal_(mean=0.0, std=output_std)
        for b in (self.attn_qkvb, self.attn_ob, self.attn_nb, self.inter_b, self.output_b, self.norm_b):
            b.data.zero_()
        self.attn_nw.data.fill_(1.0)
        self.norm_w.data.fill_(1.0)

    def _linear(self, x, w, b=None):
        if x.is_cuda and x.dtype == torch.bfloat16 and w.dtype == torch.bfloat16:
            y = linear(x, w)
            return y if b is None else y + b
        return F.linear(x, w, b)

    def _attention(self, x, attention_mask):
        B, S, H = x.shape
        heads = self.config.heads
        d = H // heads
        qkv2d = self._linear(x, self.attn_qkvw, self.attn_qkvb)
        p_drop = self.config.attn_dropout_ratio if self.training and self.config.attn_dropout_ratio > 0 else 0.0
        if (attention_mask is None and p_drop == 0.0 and qkv2d.is_cuda and qkv2d.dtype == torch.bfloat16 and S % 128 == 0 and d in (64, 128)):
            # unmasked, no attention dropout: the packed projection [B*S, (q | k | v heads) * d] is exactly what the tcgen05 flash
            # attention kernels read (csrc/attn/flash_fwd*.cu, flash_bwd.cu) -- no head transposes, no S x S matrix in HBM
            from ..flash_attention import flash_attention_packed
            ctx = flash_attention_packed(qkv2d.reshape(B * S, 3 * H), B, S, heads, heads, d, causal=False)
            return self._linear(ctx.view(B, S, H), self.attn_ow, self.attn_ob)
        qkv = qkv2d.view(B, S, 3, heads, d)
        q, k, v = (qkv[:, :, i].transpose(1, 2) for i in range(3))
        mask = None
        if attention_mask is not None:
            mask = attention_mask.to(q.dtype)
            while mask.dim() < 4:
                mask = mask.unsqueeze(1)
        p = self.config.attn_dropout_ratio if self.training and self.config.attn_dropout_ratio > 0 else 0.0
        ctx = F.scaled_dot_product_attention(q, k, v, attn_mask=mask, dropout_p=p)
        return self._linear(ctx.transpose(1, 2).reshape(B, S, H), self.attn_ow, self.attn_ob)

    def _mlp(self, x):
        if x.is_cuda and x.dtype == torch.bfloat16 and self.inter_w.dtype == torch.bfloat16:
            h = _BiasGelu.apply(linear(x, self.inter_w), self.inter_b)
        else:
            h = F.gelu(F.linear(x, self.inter_w, self.inter_b), approximate="tanh")
        return self._linear(h, self.output_w, self.output_b)

    def forward(self, hidden_states, attention_mask=None, head_mask=None, layer_head_mask=None, encoder_hidden_states=None,
                encoder_attention_mask=None, past_key_value=None, output_attentions=False, grads=None):
        self.config.is_grad_enabled = torch.is_grad_enabled()
        self.config.training = self.training
        eps = self.config.layer_norm_eps
        drop = (lambda t: F.dropout(t, self.config.hidden_dropout_ratio, True)) if (self.training and self.config.hidden_dropout_ratio > 0) else (lambda t: t)
        x = hidden_states
        if self.config.pre_layer_norm:
            a = self._attention(fused_ops.layer_norm(x, self.attn_nw, self.attn_nb, eps), attention_mask)
            x = x + drop(a)
            x = x + drop(self._mlp(fused_ops.layer_norm(x, self.norm_w, self.norm_b, eps)))
        else:
            a = self._attention(x, attention_mask)
            x = fused_ops.layer_norm(x + drop(a), self.attn_nw, self.attn_nb, eps)
            x = fused_ops.layer_norm(x + drop(self._mlp(x)), self.norm_w, self.norm_b, eps)
        return (x, ) if self.config.return_tuple else x
