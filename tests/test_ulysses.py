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
