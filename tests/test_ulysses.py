"""Ulysses DistributedAttention == full attention on the un-sharded sequence (gloo, 2 ranks), forward and backward."""
import torch

from common import run_distributed


def _body():
    import torch.nn.functional as F
    from deepspeedsyclsupport_b200 import comm as dist
    from deepspeedsyclsupport_b200.sequence.layer import DistributedAttention
    rank, P = dist.get_rank(), dist.get_world_size()
    torch.manual_seed(0)
    b, s, h, d = 2, 16, 4, 8
    q, k, v = (torch.randn(s, b, h, d, requires_grad=True) for _ in range(3))

    def local_attn(q_, k_, v_):  # [s, b, h/P, d]
        o = F.scaled_dot_product_attention(q_.permute(1, 2, 0, 3), k_.permute(1, 2, 0, 3), v_.permute(1, 2, 0, 3), is_causal=True)
        return o.permute(2, 0, 1, 3)

    ref = local_attn(q, k, v)
    ref.sum().backward()
    gq = q.grad.clone()
    sl = slice(rank * s // P, (rank + 1) * s // P)
    ql, kl, vl = (t.detach()[sl].clone().requires_grad_() for t in (q, k, v))
    attn = DistributedAttention(local_attn, dist.get_world_group())
    out = attn(ql, kl, vl)
    assert torch.allclose(out, ref[sl], atol=1e-5)
    out.sum().backward()
    assert torch.allclose(ql.grad, gq[sl], atol=1e-5)


def test_ulysses_two_ranks():
    run_distributed(_body, world_size=2)


def _packed_exchange_body():
    """sequence/fused.py on the gloo fall-back: packed QKV re-sharding round trip and agreement with a single-process
    attention over the whole sequence."""
    import torch
    from deepspeedsyclsupport_b200 import comm as dist
    from deepspeedsyclsupport_b200.sequence.fused import attn_to_sequence, qkv_to_heads
    P, rank = dist.get_world_size(), dist.get_rank()
    B, S, hq, hkv, D = 2, 8, 4, 2, 8
    torch.manual_seed(0)
    full = torch.randn(B, S, (hq + 2 * hkv) * D)           # the same tensor on every rank
    S_local = S // P
    mine = full[:, rank * S_local:(rank + 1) * S_local].reshape(B * S_local, -1).contiguous()
    heads = qkv_to_heads(mine, B, S_local, hq, hkv, D, None)  # [B * S, (hq/P + 2 hkv/P) D]
    q, k, v = full.view(B, S, hq + 2 * hkv, D).split([hq, hkv, hkv], dim=2)
    want = torch.cat([q[:, :, rank * hq // P:(rank + 1) * hq // P], k[:, :, rank * hkv // P:(rank + 1) * hkv // P],
                      v[:, :, rank * hkv // P:(rank + 1) * hkv // P]], dim=2).reshape(B * S, -1)
    assert torch.equal(heads, want)
    # the output exchange is the inverse on the q segment
    y = want[:, :hq // P * D].contiguous()
    back = attn_to_sequence(y, B, S_local, hq, D, None)
    assert torch.equal(back, q[:, rank * S_local:(rank + 1) * S_local].reshape(B * S_local, -1))


def test_packed_ulysses_exchange_gloo():
    run_distributed(_packed_exchange_body, world_size=2)
