"""DeepSpeedTransformerLayer vs. a plain BERT layer, Evoformer attention vs. the explicit formula, spatial bias-add
(CPU paths here; the CUDA paths are exercised in test_kernels_gpu.py)."""
import math

import pytest
import torch
import torch.nn.functional as F


def _bert_layer_ref(x, mask, L, pre_ln, heads, eps):
    def attn(h):
        B, S, H = h.shape
        qkv = F.linear(h, L.attn_qkvw, L.attn_qkvb).view(B, S, 3, heads, H // heads)
        q, k, v = (qkv[:, :, i].transpose(1, 2) for i in range(3))
        s = q @ k.transpose(-1, -2) / math.sqrt(H // heads)
        if mask is not None:
            s = s + mask
        return F.linear((torch.softmax(s, -1) @ v).transpose(1, 2).reshape(B, S, H), L.attn_ow, L.attn_ob)

    def mlp(h):
        return F.linear(F.gelu(F.linear(h, L.inter_w, L.inter_b), approximate="tanh"), L.output_w, L.output_b)

    ln1 = lambda t: F.layer_norm(t, (t.shape[-1], ), L.attn_nw, L.attn_nb, eps)
    ln2 = lambda t: F.layer_norm(t, (t.shape[-1], ), L.norm_w, L.norm_b, eps)
    if pre_ln:
        x = x + attn(ln1(x))
        return x + mlp(ln2(x))
    x = ln1(x + attn(x))
    return ln2(x + mlp(x))


@pytest.mark.parametrize("pre_ln", [True, False])
def test_transformer_layer_matches_bert_math(pre_ln):
    from deepspeedsyclsupport_b200.ops.transformer import DeepSpeedTransformerConfig, DeepSpeedTransformerLayer
    torch.manual_seed(0)
    cfg = DeepSpeedTransformerConfig(batch_size=2, hidden_size=32, intermediate_size=64, heads=4, attn_dropout_ratio=0.0, hidden_dropout_ratio=0.0,
                                     num_hidden_layers=2, initializer_range=0.2, pre_layer_norm=pre_ln, layer_norm_eps=1e-5)
    layer = DeepSpeedTransformerLayer(cfg)
    for b in (layer.attn_qkvb, layer.inter_b, layer.output_b, layer.attn_nb):
        b.data.normal_(std=0.1)
    x = torch.randn(2, 9, 32, requires_grad=True)
    mask = torch.zeros(2, 1, 1, 9)
    mask[1, ..., -3:] = -1e4
    y = layer(x, attention_mask=mask)
    ref = _bert_layer_ref(x, mask, layer, pre_ln, 4, 1e-5)
    assert torch.allclose(y, ref, atol=1e-5), (y - ref).abs().max()
    gy = torch.autograd.grad(y.sum(), layer.inter_w, retain_graph=True)[0]
    gr = torch.autograd.grad(ref.sum(), layer.inter_w)[0]
    assert torch.allclose(gy, gr, atol=1e-4)


def test_evoformer_attention_with_bias_grads():
    from deepspeedsyclsupport_b200.ops.deepspeed4science import DS4Sci_EvoformerAttention
    torch.manual_seed(0)
    B, N, L, H, D = 1, 3, 20, 2, 8
    q, k, v = (torch.randn(B, N, L, H, D, requires_grad=True) for _ in range(3))
    b1 = torch.randn(B, N, 1, 1, L, requires_grad=True)
    b2 = torch.randn(B, 1, H, L, L, requires_grad=True)
    out = DS4Sci_EvoformerAttention(q, k, v, [b1, b2])
    qh, kh, vh = (t.transpose(-2, -3) for t in (q, k, v))
    ref = (torch.softmax(qh @ kh.transpose(-1, -2) / math.sqrt(D) + b1 + b2, -1) @ vh).transpose(-2, -3)
    assert out.shape == q.shape and torch.allclose(out, ref, atol=1e-5)
    g = torch.randn_like(out)
    got = torch.autograd.grad(out, (q, k, v, b1, b2), g)
    want = torch.autograd.grad(ref, (q, k, v, b1, b2), g)
    for a, b in zip(got, want):
        assert a.shape == b.shape and torch.allclose(a, b, atol=1e-4)
    assert torch.allclose(DS4Sci_EvoformerAttention(q, k, v, []), (torch.softmax(qh @ kh.transpose(-1, -2) / math.sqrt(D), -1) @ vh).transpose(-2, -3), atol=1e-5)


def test_nhwc_bias_add_cpu():
    from deepspeedsyclsupport_b200.ops.spatial import nhwc_bias_add, nhwc_bias_add_add, nhwc_bias_add_bias_add
    x = torch.randn(2, 8, 4, 4).contiguous(memory_format=torch.channels_last)
    o = torch.randn(2, 8, 4, 4).contiguous(memory_format=torch.channels_last)
    b, ob = torch.randn(8), torch.randn(8)
    assert torch.allclose(nhwc_bias_add(x, b), x + b.view(1, -1, 1, 1))
    assert torch.allclose(nhwc_bias_add_add(x, b, o), x + b.view(1, -1, 1, 1) + o)
    assert torch.allclose(nhwc_bias_add_bias_add(x, b, o, ob), x + b.view(1, -1, 1, 1) + o + ob.view(1, -1, 1, 1))
