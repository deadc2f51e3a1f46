"""Block-sparse attention vs. a dense masked fp32 reference (reference tests: tests/unit/ops/sparse_attention)."""
import pytest
import torch


def _dense_ref(q, k, v, layout, block, key_padding_mask=None, attn_mask=None, kp_mode="add", am_mode="mul"):
    B, H, S, D = q.shape
    mask = layout.bool().repeat_interleave(block, 1).repeat_interleave(block, 2)  # [H, S, S]
    s = (q.float() @ k.float().transpose(-1, -2)) * D**-0.5
    if attn_mask is not None:
        s = s + attn_mask.float() if am_mode == "add" else s.masked_fill(attn_mask == 0, float("-inf"))
    if key_padding_mask is not None:
        kp = key_padding_mask.float().view(B, 1, 1, S)
        s = s + kp if kp_mode == "add" else s.masked_fill(kp == 0, float("-inf"))
    s = s.masked_fill(~mask.unsqueeze(0), float("-inf"))
    return torch.softmax(s, -1) @ v.float()


@pytest.mark.parametrize("cfg_name,kw", [
    ("FixedSparsityConfig", dict(num_local_blocks=2, num_global_blocks=1, attention="unidirectional")),
    ("BigBirdSparsityConfig", dict(num_random_blocks=1, num_sliding_window_blocks=3, num_global_blocks=1)),
    ("BSLongformerSparsityConfig", dict(num_sliding_window_blocks=3, global_block_indices=[0])),
    ("LocalSlidingWindowSparsityConfig", dict(num_sliding_window_blocks=3)),
    ("DenseSparsityConfig", dict()),
])
def test_sparse_self_attention_matches_dense(cfg_name, kw):
    from deepspeedsyclsupport_b200.ops import sparse_attention as SA
    torch.manual_seed(0)
    B, H, S, D, blk = 2, 4, 128, 32, 16
    cfg = getattr(SA, cfg_name)(num_heads=H, block=blk, **kw)
    attn = SA.SparseSelfAttention(cfg, max_seq_length=256)
    q, k, v = (torch.randn(B, H, S, D, requires_grad=True) for _ in range(3))
    layout = attn.get_layout(S)
    kp = torch.zeros(B, S)
    kp[1, -20:] = -10000.0
    am = torch.tril(torch.ones(S, S)) if "unidirectional" in str(kw) else None
    out = attn(q, k, v, key_padding_mask=kp, attn_mask=am)
    ref = _dense_ref(q, k, v, layout, blk, key_padding_mask=kp, attn_mask=am)
    assert torch.allclose(out, ref, atol=1e-4), (out - ref).abs().max()
    g = torch.randn_like(out)
    gq, gk, gv = torch.autograd.grad(out, (q, k, v), g)
    rq, rk, rv = torch.autograd.grad(ref, (q, k, v), g)
    for a, b in ((gq, rq), (gk, rk), (gv, rv)):
        assert torch.allclose(a, b, atol=1e-4)


def test_block_sparse_matmul_modes():
    from deepspeedsyclsupport_b200.ops.sparse_attention import MatMul
    torch.manual_seed(0)
    H, nb, blk, D, B = 2, 4, 8, 16, 3
    layout = (torch.rand(H, nb, nb) > 0.5).long()
    layout[:, 0, 0] = 1
    mask = layout.bool().repeat_interleave(blk, 1).repeat_interleave(blk, 2)
    a, b = torch.randn(B, H, nb * blk, D), torch.randn(B, H, nb * blk, D)
    sdd = MatMul(layout, blk, "sdd", trans_b=True)(a, b)
    dense = (a @ b.transpose(-1, -2)) * mask
    h, r, c = layout.nonzero(as_tuple=True)
    got = torch.zeros_like(dense).view(B, H, nb, blk, nb, blk)
    got[:, h, r, :, c, :] = sdd.transpose(0, 1)
    assert torch.allclose(got.view_as(dense), dense, atol=1e-5)
    dsd = MatMul(layout, blk, "dsd")(sdd, b)
    assert torch.allclose(dsd, dense @ b, atol=1e-4)
    dds = MatMul(layout, blk, "dds")(a.transpose(-1, -2), sdd)
    assert torch.allclose(dds, a.transpose(-1, -2) @ dense, atol=1e-4)


def test_pad_to_block_size():
    from deepspeedsyclsupport_b200.ops.sparse_attention import SparseAttentionUtils
    ids = torch.ones(2, 30, dtype=torch.long)
    pad_len, ids2, am, tt, pos, emb = SparseAttentionUtils.pad_to_block_size(16, ids, torch.ones(2, 30), torch.zeros(2, 30, dtype=torch.long), None, None, 0, None)
    assert pad_len == 2 and ids2.shape == (2, 32) and am[:, -2:].sum() == 0
    assert SparseAttentionUtils.unpad_sequence_output(pad_len, torch.zeros(2, 32, 4)).shape == (2, 30, 4)
