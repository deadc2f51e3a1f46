"""Numerics of every CUDA kernel against a plain PyTorch fp32 reference of the same op (GPU only)."""
import math

import pytest
import torch

pytestmark = pytest.mark.gpu


def _dev():
    return torch.device("cuda")


@pytest.mark.parametrize("layout", [(False, False), (False, True), (True, True), (True, False)])
@pytest.mark.parametrize("shape", [(256, 512, 128), (1000, 1096, 328), (2048, 4096, 1024), (192, 64, 128)])
def test_gemm_bf16(layout, shape):
    from deepspeedsyclsupport_b200.ops.gemm import gemm
    a_mn, b_mn = layout
    M, N, K = shape
    torch.manual_seed(0)
    a = torch.randn((K, M) if a_mn else (M, K), device=_dev(), dtype=torch.bfloat16)
    b = torch.randn((K, N) if b_mn else (N, K), device=_dev(), dtype=torch.bfloat16)
    ref = (a.t() if a_mn else a).float() @ (b if b_mn else b.t()).float()
    c = gemm(a, b, a_mn=a_mn, b_mn=b_mn)
    assert (c.float() - ref).abs().max() <= 1e-2 * ref.abs().max()
    c32 = gemm(a, b, a_mn=a_mn, b_mn=b_mn, out_dtype=torch.float32)
    assert (c32 - ref).abs().max() <= 1e-4 * ref.abs().max()
    acc = torch.randn(M, N, device=_dev())
    want = acc + ref
    gemm(a, b, a_mn=a_mn, b_mn=b_mn, out=acc, accumulate=True)
    assert (acc - want).abs().max() <= 1e-4 * want.abs().max()


def test_linear_autograd():
    from deepspeedsyclsupport_b200.ops.gemm import linear
    torch.manual_seed(0)
    x = torch.randn(4, 96, 256, device=_dev(), dtype=torch.bfloat16, requires_grad=True)
    w = torch.randn(384, 256, device=_dev(), dtype=torch.bfloat16, requires_grad=True)
    y = linear(x, w)
    dy = torch.randn_like(y)
    y.backward(dy)
    xr, wr = x.detach().float().requires_grad_(), w.detach().float().requires_grad_()
    yr = torch.nn.functional.linear(xr, wr)
    yr.backward(dy.float())
    assert (y.float() - yr).abs().max() < 2e-2 * yr.abs().max()
    assert (x.grad.float() - xr.grad).abs().max() < 2e-2 * xr.grad.abs().max()
    assert (w.grad.float() - wr.grad).abs().max() < 2e-2 * wr.grad.abs().max()


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16, torch.float32])
@pytest.mark.parametrize("hidden", [256, 4096, 5120, 8192])
def test_rms_norm_fwd_bwd(dtype, hidden):
    from deepspeedsyclsupport_b200.ops import fused_ops as FO
    torch.manual_seed(0)
    rows = 300
    x = torch.randn(rows, hidden, device=_dev(), dtype=dtype, requires_grad=True)
    res = torch.randn(rows, hidden, device=_dev(), dtype=dtype, requires_grad=True)
    w = (1 + 0.1 * torch.randn(hidden, device=_dev())).to(dtype).requires_grad_()
    y, h = FO.rms_norm(x, w, 1e-5, residual=res)
    dy, dh = torch.randn_like(y), torch.randn_like(h)
    torch.autograd.backward([y, h], [dy, dh])
    xr, rr, wr = (t.detach().float().requires_grad_() for t in (x, res, w))
    hr = xr + rr
    yr = hr * torch.rsqrt(hr.pow(2).mean(-1, keepdim=True) + 1e-5) * wr
    torch.autograd.backward([yr, hr], [dy.float(), dh.float()])
    tol = 2e-2 if dtype != torch.float32 else 1e-4
    assert (y.float() - yr).abs().max() < tol * max(1.0, yr.abs().max().item())
    assert (x.grad.float() - xr.grad).abs().max() < tol * max(1.0, xr.grad.abs().max().item())
    assert (res.grad.float() - rr.grad).abs().max() < tol * max(1.0, rr.grad.abs().max().item())
    assert (w.grad.float() - wr.grad).abs().max() < 3 * tol * max(1.0, wr.grad.abs().max().item())


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float32])
def test_layer_norm_fwd_bwd(dtype):
    from deepspeedsyclsupport_b200.ops import fused_ops as FO
    torch.manual_seed(0)
    rows, hidden = 257, 1024
    x = torch.randn(rows, hidden, device=_dev(), dtype=dtype, requires_grad=True)
    w = (1 + 0.1 * torch.randn(hidden, device=_dev())).to(dtype).requires_grad_()
    b = (0.1 * torch.randn(hidden, device=_dev())).to(dtype).requires_grad_()
    y = FO.layer_norm(x, w, b, 1e-5)
    dy = torch.randn_like(y)
    y.backward(dy)
    xr, wr, br = (t.detach().float().requires_grad_() for t in (x, w, b))
    yr = torch.nn.functional.layer_norm(xr, (hidden, ), wr, br, 1e-5)
    yr.backward(dy.float())
    tol = 2e-2 if dtype != torch.float32 else 1e-4
    assert (y.float() - yr).abs().max() < tol * max(1.0, yr.abs().max().item())
    assert (x.grad.float() - xr.grad).abs().max() < tol * max(1.0, xr.grad.abs().max().item())
    assert (w.grad.float() - wr.grad).abs().max() < 3 * tol * max(1.0, wr.grad.abs().max().item())
    assert (b.grad.float() - br.grad).abs().max() < 3 * tol * max(1.0, br.grad.abs().max().item())


def test_rope_fwd_bwd():
    from deepspeedsyclsupport_b200.ops import fused_ops as FO
    torch.manual_seed(0)
    B, S, hq, hkv, d = 2, 128, 4, 2, 128
    inv = FO.rope_inv_freq(d, device=_dev())
    qkv = torch.randn(B * S, (hq + 2 * hkv) * d, device=_dev(), dtype=torch.bfloat16)
    pos = (torch.arange(B * S, device=_dev()) % S)
    ref = FO.rope_ref(qkv.float(), inv, pos, 0, hq + hkv, d)
    x = qkv.clone().requires_grad_()
    y = FO.rope_(x * 1.0, inv, S, 0, hq + hkv, d)
    assert (y.float() - ref).abs().max() < 3e-2
    dy = torch.randn_like(y)
    y.backward(dy)
    ref_b = FO.rope_ref(dy.float(), inv, pos, 0, hq + hkv, d, backward=True)
    assert (x.grad.float() - ref_b).abs().max() < 3e-2


@pytest.mark.parametrize("act", ["silu", "gelu", "relu"])
def test_gated_act(act):
    from deepspeedsyclsupport_b200.ops import fused_ops as FO
    torch.manual_seed(0)
    x = torch.randn(333, 2 * 1376, device=_dev(), dtype=torch.bfloat16, requires_grad=True)
    y = FO.gated_act(x, act)
    dy = torch.randn_like(y)
    y.backward(dy)
    xr = x.detach().float().requires_grad_()
    yr = FO.gated_act_ref(xr, act)
    yr.backward(dy.float())
    assert (y.float() - yr).abs().max() < 3e-2 * max(1.0, yr.abs().max().item())
    assert (x.grad.float() - xr.grad).abs().max() < 3e-2 * max(1.0, xr.grad.abs().max().item())


def test_cross_entropy():
    from deepspeedsyclsupport_b200.ops import fused_ops as FO
    torch.manual_seed(0)
    T, V = 513, 32000
    logits = (3 * torch.randn(T, V, device=_dev())).bfloat16()
    labels = torch.randint(0, V, (T, ), device=_dev())
    labels[::7] = -100
    lr = logits.float().clone().requires_grad_()
    ref = torch.nn.functional.cross_entropy(lr, labels, ignore_index=-100)
    ref.backward()
    l = logits.clone().requires_grad_()
    loss = FO.cross_entropy(l * 1.0, labels)
    loss.backward()
    assert abs(loss.item() - ref.item()) < 2e-3 * abs(ref.item())
    assert (l.grad.float() - lr.grad).abs().max() < 2e-2 * lr.grad.abs().max() + 1e-6


@pytest.mark.parametrize("gdt", [torch.float32, torch.bfloat16])
def test_fused_adam_flat(gdt):
    from deepspeedsyclsupport_b200.ops.flat_optim import adam_ref, adam_step
    torch.manual_seed(0)
    n = 1_000_003
    p = torch.randn(n, device=_dev())
    m, v = torch.zeros_like(p), torch.zeros_like(p)
    p2, m2, v2 = p.clone(), m.clone(), v.clone()
    lp = torch.empty(n, device=_dev(), dtype=torch.bfloat16)
    inv = torch.tensor([0.5], device=_dev())
    for step in range(1, 4):
        g = torch.randn(n, device=_dev()).to(gdt)
        adam_step(p, m, v, g, lr=1e-2, weight_decay=0.1, step=step, lp=lp, inv_scale=inv)
        adam_ref(p2, m2, v2, g, 1e-2, 0.9, 0.999, 1e-8, 0.1, True, True, step, grad_scale=0.5)
    assert (p - p2).abs().max() < 1e-5
    assert (m - m2).abs().max() < 1e-5 and (v - v2).abs().max() < 1e-5
    assert (lp.float() - p2).abs().max() < 1e-2 * p2.abs().max()


def test_fused_adam_optimizer_class_and_lion_lamb():
    from deepspeedsyclsupport_b200.ops.adam import FusedAdam
    from deepspeedsyclsupport_b200.ops.lamb import FusedLamb
    from deepspeedsyclsupport_b200.ops.lion import FusedLion
    torch.manual_seed(0)
    ws = [torch.randn(257, 33, device=_dev(), requires_grad=True), torch.randn(70000, device=_dev(), requires_grad=True)]
    wr = [w.detach().clone().requires_grad_() for w in ws]
    o, r = FusedAdam(ws, lr=1e-2, weight_decay=0.05), torch.optim.AdamW(wr, lr=1e-2, weight_decay=0.05)
    for _ in range(3):
        for a, b in zip(ws, wr):
            g = torch.randn_like(a)
            a.grad, b.grad = g.clone(), g.clone()
        o.step()
        r.step()
    for a, b in zip(ws, wr):
        assert (a - b).abs().max() < 1e-5
    # Lion: sign update
    w = torch.randn(5000, device=_dev(), requires_grad=True)
    w0 = w.detach().clone()
    g = torch.randn_like(w)
    w.grad = g.clone()
    FusedLion([w], lr=1e-2, betas=(0.9, 0.99), weight_decay=0.0).step()
    assert torch.allclose(w.detach(), w0 - 1e-2 * torch.sign(0.1 * g), atol=1e-6)
    # LAMB: trust ratio
    w = torch.randn(40000, device=_dev(), requires_grad=True)
    w0 = w.detach().clone()
    g = torch.randn_like(w)
    w.grad = g.clone()
    opt = FusedLamb([w], lr=1e-2, weight_decay=0.01, bias_correction=True)
    opt.step()
    mm, vv = 0.1 * g, 0.001 * g * g
    upd = (mm / 0.1) / ((vv / 0.001).sqrt() + 1e-8) + 0.01 * w0
    ratio = min(max((w0.norm() / upd.norm()).item(), 0.01), 10.0)
    assert torch.allclose(w.detach(), w0 - 1e-2 * ratio * upd, atol=1e-5)
    assert abs(opt.get_lamb_coeffs()[0] - ratio) < 1e-4


@pytest.mark.parametrize("cfg", [(2, 256, 2, 2, 128, True), (2, 512, 4, 2, 128, True), (1, 384, 2, 2, 64, True), (2, 256, 2, 1, 128, False)])
def test_flash_attention_fwd_bwd(cfg):
    from deepspeedsyclsupport_b200.ops.attention import attention, attention_ref
    B, S, hq, hkv, d, causal = cfg
    torch.manual_seed(0)
    qkv = torch.randn(B * S, (hq + 2 * hkv) * d, device=_dev(), dtype=torch.bfloat16, requires_grad=True)
    out = attention(qkv, B, S, hq, hkv, d, causal=causal)
    dout = torch.randn_like(out)
    out.backward(dout)
    qf = qkv.detach().float().requires_grad_()
    ref = attention_ref(qf, B, S, hq, hkv, d, causal)
    ref.backward(dout.float())
    assert (out.float() - ref).abs().max() < 2e-2
    assert (qkv.grad.float() - qf.grad).abs().max() < 2e-2 * max(1.0, qf.grad.abs().max().item())


def test_flash_attention_forward_lazy_rescale_and_long_sequence():
    """Scores that keep growing along the key axis force the forward's lazy O rescale (row max jumps by more than 2^8 between
    key blocks); 2048 tokens exercise the two-Q-tile ping-pong over many key blocks.  Also checks LSE."""
    from deepspeedsyclsupport_b200.ops.attention import attention_ref
    from deepspeedsyclsupport_b200.ops.flash_attention import flash_fwd
    B, S, hq, hkv, d = 1, 2048, 2, 1, 128
    torch.manual_seed(0)
    qkv = torch.randn(B * S, (hq + 2 * hkv) * d, device=_dev(), dtype=torch.float32)
    ramp = torch.linspace(0.05, 12.0, S, device=_dev()).repeat(B)[:, None]
    qkv[:, hq * d:(hq + hkv) * d] *= ramp  # keys grow with position: later blocks dominate the softmax
    qkv = qkv.bfloat16()
    for causal in (True, False):
        out, lse = flash_fwd(qkv, B, S, hq, hkv, d, causal)
        ref = attention_ref(qkv.float(), B, S, hq, hkv, d, causal)
        assert torch.isfinite(out.float()).all() and (out.float() - ref).abs().max() < 3e-2
        q = qkv.float().view(B, S, hq + 2 * hkv, d)
        s = torch.einsum("bihd,bjd->bhij", q[:, :, :hq], q[:, :, hq]) / math.sqrt(d)
        if causal:
            s = s.masked_fill(torch.ones(S, S, device=_dev(), dtype=torch.bool).triu(1), float("-inf"))
        assert (lse - torch.logsumexp(s, -1)).abs().max() < 2e-2 * max(1.0, s.abs().max().item() / 50)


def test_p2p_collectives_emulated_ranks():
    """All 'ranks' live on one GPU (one buffer each); each rank's kernel is launched in turn (no cross-kernel waits)."""
    from deepspeedsyclsupport_b200.parallel.symm import EmulatedPeers
    torch.manual_seed(0)
    W, n = 4, 8 * 1000
    em = EmulatedPeers(W, nbytes=W * n * 4)
    # all-gather: rank r's shard = n bf16 at offset 0 of its buffer
    shards = [torch.randn(n, device=_dev()).bfloat16() for _ in range(W)]
    for r in range(W):
        em.view(0, (n, ), torch.bfloat16, rank=r).copy_(shards[r])
    for r in range(W):
        out = torch.empty(W * n, device=_dev(), dtype=torch.bfloat16)
        em.all_gather(out, 0, n * 2, rank=r)
        assert torch.equal(out, torch.cat(shards))
    # all-to-all: rank r's send buffer holds W chunks; rank q must receive chunk q of every rank
    sends = [torch.randn(W * n, device=_dev()).bfloat16() for _ in range(W)]
    for r in range(W):
        em.view(0, (W * n, ), torch.bfloat16, rank=r).copy_(sends[r])
    for q in range(W):
        out = torch.empty(W * n, device=_dev(), dtype=torch.bfloat16)
        em.all_to_all(out, 0, n * 2, rank=q)
        assert torch.equal(out, torch.cat([sends[r][q * n:(q + 1) * n] for r in range(W)]))
    # reduce-scatter (+scale, +norm): every rank holds W*n bf16 gradients
    grads = [torch.randn(W * n, device=_dev()).bfloat16() for _ in range(W)]
    for r in range(W):
        em.view(0, (W * n, ), torch.bfloat16, rank=r).copy_(grads[r])
    total = sum(g.float() for g in grads) / W
    for r in range(W):
        dst = torch.zeros(n, device=_dev())
        sq = torch.zeros(1, device=_dev())
        em.reduce_scatter(dst, 0, torch.bfloat16, n, scale=1.0 / W, sumsq=sq, rank=r)
        want = total[r * n:(r + 1) * n]
        assert (dst - want).abs().max() < 1e-5
        assert abs(sq.item() - want.square().sum().item()) < 1e-2 * want.square().sum().item()
        dst16 = torch.zeros(n, device=_dev(), dtype=torch.bfloat16)
        em.reduce_scatter(dst16, 0, torch.bfloat16, n, scale=1.0 / W, rank=r)
        assert (dst16.float() - want).abs().max() < 2e-2
    # fused reduce-scatter + AdamW
    from deepspeedsyclsupport_b200.ops.flat_optim import adam_ref
    for r in range(W):
        p = torch.randn(n, device=_dev())
        m, v = torch.zeros(n, device=_dev()), torch.zeros(n, device=_dev())
        p2, m2, v2 = p.clone(), m.clone(), v.clone()
        lp = torch.empty(n, device=_dev(), dtype=torch.bfloat16)
        em.rs_adam(0, p, m, v, lp, n, scale=1.0 / W, lr=1e-2, beta1=0.9, beta2=0.95, eps=1e-8, weight_decay=0.1, step=1, rank=r)
        adam_ref(p2, m2, v2, total[r * n:(r + 1) * n], 1e-2, 0.9, 0.95, 1e-8, 0.1, True, True, 1)
        assert (p - p2).abs().max() < 1e-5 and (lp.float() - p2).abs().max() < 2e-2
    # one-shot all-reduce
    for r in range(W):
        dst = torch.empty(W * n, device=_dev(), dtype=torch.bfloat16)
        em.all_reduce_bf16(dst, 0, W * n, rank=r)
        assert (dst.float() - total * W).abs().max() < 5e-2


@pytest.mark.parametrize("shape", [(1024, 1376, 512), (2048, 5120, 1024)])  # single-CTA core / CTA-pair core
@pytest.mark.parametrize("world", [1, 4])
def test_ag_gemm_emulated_ranks(world, shape):
    """All-gather fused into the GEMM: W is sharded over `world` flat shards (all on this GPU); the kernel pulls the
    shards through the peer-pointer table while the tcgen05 tiles run."""
    from deepspeedsyclsupport_b200.parallel.fused import ag_gemm
    torch.manual_seed(0)
    M, N, K = shape
    w_off = 64 * 5  # W starts in the middle of the unit
    unit = w_off + N * K + 128
    S = -(-unit // (world * 64)) * 64
    flat = torch.randn(world * S, device=_dev()).bfloat16()
    W = flat[w_off:w_off + N * K].view(N, K)
    shards = [flat[r * S:(r + 1) * S].clone() for r in range(world)]
    x = torch.randn(M, K, device=_dev(), dtype=torch.bfloat16)
    w_local = torch.zeros(N, K, device=_dev(), dtype=torch.bfloat16)
    out = torch.empty(M, N, device=_dev(), dtype=torch.bfloat16)
    flags = torch.empty(64, device=_dev(), dtype=torch.int32)
    for _ in range(2):  # second call: stale flags / buffers must not matter
        w_local.zero_()
        ag_gemm(x, w_local, out, [s.data_ptr() for s in shards], w_off, S, flags, N, K)
        torch.cuda.synchronize()
        assert torch.equal(w_local, W)
        ref = x.float() @ W.float().t()
        assert (out.float() - ref).abs().max() <= 1e-2 * ref.abs().max()


def test_fp8_block_quantizer_and_reduce():
    from deepspeedsyclsupport_b200.ops import quantizer as Q
    torch.manual_seed(0)
    x = (torch.randn(128 * 1000, device=_dev()) * 3).bfloat16()
    q, s = Q.quantize_fp8_block(x)
    qr, sr = Q.quantize_fp8_block_ref(x)
    assert torch.allclose(s, sr, rtol=1e-6)
    d = Q.dequantize_fp8_block(q, s)
    dr = Q.dequantize_fp8_block_ref(qr, sr)
    assert (d.float() - dr.float()).abs().max() <= 0.07 * x.float().abs().max()  # at most one fp8 ulp apart
    assert (d.float() - x.float()).abs().max() <= 0.07 * x.float().abs().max()
    parts = 4
    xs = torch.randn(parts, 128 * 50, device=_dev()).bfloat16()
    q, s = Q.quantize_fp8_block(xs.view(-1))
    out = Q.dequant_reduce_fp8_block(q, s, parts, dtype=torch.float32)
    ref = Q.dequantize_fp8_block(q, s, dtype=torch.float32).view(parts, -1).sum(0)
    assert (out - ref).abs().max() < 1e-4


@pytest.mark.parametrize("bits,sym", [(8, True), (8, False), (4, True), (4, False)])
def test_int_quantizer(bits, sym):
    from deepspeedsyclsupport_b200.ops import quantizer as Q
    torch.manual_seed(0)
    groups, gs = 64, 512
    x = torch.randn(groups * gs, device=_dev(), dtype=torch.float16)
    q, params = Q.int_quantize(x, groups, bits, sym)
    d = Q.int_dequantize(q, params, groups, bits, sym, torch.float32)
    qr, pr = Q.int_quantize_ref(x, groups, bits, sym)
    dr = Q.int_dequantize_ref(qr, pr).view(-1)
    assert torch.allclose(params, pr, rtol=1e-5, atol=1e-7)
    step = params[:, 0].max().item()
    assert (d - dr).abs().max() <= step * 1.01  # rounding ties may differ by one level
    assert (d - x.float()).abs().max() <= step * 1.01


def test_fake_quantize():
    from deepspeedsyclsupport_b200.ops import quantizer as Q
    torch.manual_seed(0)
    x = torch.randn(32, 256, device=_dev(), dtype=torch.float16)
    for sym in (True, False):
        y = Q.fake_quantize_(x.clone(), 32, 8, symmetric=sym)
        yr = Q.fake_quantize_(x.clone().cpu(), 32, 8, symmetric=sym)
        step = (x.float().abs().amax(1) * 2 / 256).max().item()
        assert (y.cpu().float() - yr.float()).abs().max() <= step * 1.01
    ys = Q.fake_quantize_(x.clone(), 32, 4, symmetric=True, stochastic=True, seed=7)
    step4 = (x.float().abs().amax(1) * 2 / 16).max().item()
    assert (ys.float() - x.float()).abs().max() <= step4 * 1.01


def test_ragged_kernels_vs_reference():
    """kv append + RoPE, paged attention (decode + chunked prefill), embed, logits gather, top-k gating, scatter/gather."""
    from deepspeedsyclsupport_b200.inference.v2 import kernels as K
    from deepspeedsyclsupport_b200.ops.fused_ops import rope_inv_freq
    torch.manual_seed(0)
    dev = _dev()
    hq, hkv, d, bs, nblocks = 8, 2, 128, 16, 64
    inv = rope_inv_freq(d, device=dev)
    # two sequences: seq0 has 37 cached tokens + 5 new, seq1 is new with 20 tokens
    seen, ntok = [37, 0], [5, 20]
    table = torch.zeros(2, 8, dtype=torch.int32, device=dev)
    table[0, :3] = torch.tensor([5, 9, 2])
    table[1, :2] = torch.tensor([7, 11])
    cache = torch.zeros(nblocks, 2, bs, hkv, d, dtype=torch.bfloat16, device=dev)
    cache_ref = cache.cpu().clone()
    # history of seq 0
    hist = torch.randn(37, hq + 2 * hkv, d, device=dev).bfloat16()
    pos_h = torch.arange(37, dtype=torch.int32, device=dev)
    seq_h = torch.zeros(37, dtype=torch.int32, device=dev)
    K.kv_rotary_append(hist.view(37, -1).clone(), cache, seq_h, pos_h, table, inv, hq, hkv, d)
    K.kv_rotary_append(hist.view(37, -1).cpu().clone(), cache_ref, seq_h.cpu(), pos_h.cpu(), table.cpu(), inv.cpu(), hq, hkv, d)
    T = sum(ntok)
    qkv = torch.randn(T, (hq + 2 * hkv) * d, device=dev).bfloat16()
    tok_seq = torch.tensor([0] * 5 + [1] * 20, dtype=torch.int32, device=dev)
    tok_pos = torch.tensor(list(range(37, 42)) + list(range(20)), dtype=torch.int32, device=dev)
    q_gpu = qkv.clone()
    K.kv_rotary_append(q_gpu, cache, tok_seq, tok_pos, table, inv, hq, hkv, d)
    q_cpu = qkv.cpu().clone()
    K.kv_rotary_append(q_cpu, cache_ref, tok_seq.cpu(), tok_pos.cpu(), table.cpu(), inv.cpu(), hq, hkv, d)
    assert (q_gpu.cpu().float() - q_cpu.float()).abs().max() < 3e-2
    assert (cache.cpu().float() - cache_ref.float()).abs().max() < 3e-2
    out = K.paged_attention(q_gpu, cache, tok_seq, tok_pos, table, hq, hkv, d)
    ref = K.paged_attention_ref(q_gpu.cpu(), cache.cpu(), tok_seq.cpu(), tok_pos.cpu(), table.cpu(), hq, hkv, d, 1.0 / d ** 0.5)
    assert (out.cpu().float() - ref.float()).abs().max() < 2e-2
    # embed / logits gather
    wte = torch.randn(100, 256, device=dev).bfloat16()
    ids = torch.randint(0, 100, (T, ), device=dev)
    assert torch.equal(K.ragged_embed(ids, wte), wte[ids])
    last = torch.tensor([4, 24], dtype=torch.int32, device=dev)
    hid = torch.randn(T, 256, device=dev).bfloat16()
    assert torch.equal(K.logits_gather(hid, last), hid[last.long()])
    # MoE routing
    E, k = 8, 2
    logits = torch.randn(T, E, device=dev)
    counts, scores, asg, off = K.top_k_gating(logits, k)
    c2, s2, a2, o2 = K.top_k_gating(logits.cpu(), k)
    assert torch.equal(counts.cpu(), c2) and torch.equal(asg.cpu(), a2)
    assert torch.allclose(scores.cpu(), s2, atol=1e-5)
    for e in range(E):  # offsets are a permutation of 0..count-1 inside every expert
        sel = off[asg == e].sort().values.cpu()
        assert torch.equal(sel, torch.arange(sel.numel(), dtype=torch.int32))
    x = torch.randn(T, 256, device=dev).bfloat16()
    moe_in, cumsum, slots = K.moe_scatter(x, counts, asg, off)
    assert torch.equal(cumsum.cpu(), counts.cpu().long().cumsum(0))
    for j in range(k):
        assert torch.equal(moe_in[slots[:, j].long()], x)
    y = K.moe_gather(moe_in, scores, slots)
    assert (y.float() - x.float() * scores.sum(1, keepdim=True)).abs().max() < 3e-2


def test_fastgen_engine_gpu_matches_dense():
    import sys, os
    sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
    from test_inference_v2 import _dense_llama_logits, _hf_llama_state
    from deepspeedsyclsupport_b200.inference.v2 import RaggedInferenceEngineConfig, SplitFuseScheduler, build_engine_from_state_dict
    cfg = dict(vocab_size=512, hidden_size=512, intermediate_size=1024, num_hidden_layers=2, num_attention_heads=4, num_key_value_heads=2,
               rms_norm_eps=1e-5)
    sd = _hf_llama_state(cfg)
    ecfg = RaggedInferenceEngineConfig(state_manager={"max_ragged_batch_size": 512, "max_ragged_sequence_count": 8, "max_context": 1024,
                                                      "memory_config": {"mode": "allocate", "size": 64}})
    eng = build_engine_from_state_dict("llama", cfg, sd, ecfg)
    torch.manual_seed(1)
    a = torch.randint(0, 512, (300, ))
    ref = _dense_llama_logits({k: v.bfloat16().float() for k, v in sd.items()}, cfg, a)
    out = eng.put([1], [a[:200]])
    assert (out[0].float().cpu() - ref[199]).abs().max() < 0.06 * ref[199].abs().max() + 0.02
    out = eng.put([1], [a[200:]])
    assert (out[0].float().cpu() - ref[299]).abs().max() < 0.06 * ref[299].abs().max() + 0.02
    assert out.dtype == torch.bfloat16 and out.is_cuda


def test_fastgen_decode_steps_replay_a_cuda_graph():
    """Pure decode steps are captured into a CUDA graph per batch size after two eager steps (RaggedTransformer.forward_graphed);
    every step -- eager, the capturing one, replays, and replays after the contexts grew into new KV blocks -- must match the
    dense fp32 forward of the same token history."""
    import sys, os
    sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
    from test_inference_v2 import _dense_llama_logits, _hf_llama_state
    from deepspeedsyclsupport_b200.inference.v2 import RaggedInferenceEngineConfig, build_engine_from_state_dict
    cfg = dict(vocab_size=512, hidden_size=512, intermediate_size=1024, num_hidden_layers=2, num_attention_heads=4, num_key_value_heads=2,
               rms_norm_eps=1e-5)
    sd = _hf_llama_state(cfg)
    sd32 = {k: v.bfloat16().float() for k, v in sd.items()}
    ecfg = RaggedInferenceEngineConfig(state_manager={"max_ragged_batch_size": 512, "max_ragged_sequence_count": 8, "max_context": 1024,
                                                      "memory_config": {"mode": "allocate", "size": 64}})
    eng = build_engine_from_state_dict("llama", cfg, sd, ecfg)
    torch.manual_seed(2)
    hist = {1: torch.randint(0, 512, (120, )), 2: torch.randint(0, 512, (250, )), 3: torch.randint(0, 512, (7, ))}
    eng.put(list(hist), [hist[u] for u in hist])
    for step in range(12):  # sequence 1 crosses the 128-token block boundary at step 8, sequence 2 at step 6
        new = {u: torch.randint(0, 512, (1, )) for u in hist}
        out = eng.put(list(hist), [new[u] for u in hist]).float().cpu()
        for k, u in enumerate(hist):
            hist[u] = torch.cat([hist[u], new[u]])
            ref = _dense_llama_logits(sd32, cfg, hist[u])[-1]
            assert (out[k] - ref).abs().max() < 0.06 * ref.abs().max() + 0.02, (step, u, (out[k] - ref).abs().max().item())
    model = eng._model
    assert 3 in model._graphs and model._graph_warm[3] == 2, "decode steps of 3 sequences were not captured"
    # a different batch size gets its own warm-up and graph; the first one keeps working afterwards
    eng.flush(3)
    del hist[3]
    for step in range(4):
        new = {u: torch.randint(0, 512, (1, )) for u in hist}
        out = eng.put(list(hist), [new[u] for u in hist]).float().cpu()
        for k, u in enumerate(hist):
            hist[u] = torch.cat([hist[u], new[u]])
            ref = _dense_llama_logits(sd32, cfg, hist[u])[-1]
            assert (out[k] - ref).abs().max() < 0.06 * ref.abs().max() + 0.02, ("two sequences", step, u)
    assert 2 in model._graphs


@pytest.mark.parametrize("n", [8 * 1024, 8 * 100_003])
def test_onebit_compress_decompress(n):
    """fused sign-pack + error feedback and the unpack kernels vs. the torch formulas (runtime/comm/compressed.py CPU path)."""
    from deepspeedsyclsupport_b200.runtime.comm.compressed import onebit_compress, onebit_decompress
    torch.manual_seed(0)
    x = torch.randn(n, device=_dev())
    err = torch.randn(n, device=_dev()) * 0.1
    x_cpu, err_cpu = x.cpu().clone(), err.cpu().clone()
    packed, scale = onebit_compress(x, err)
    packed_ref, scale_ref = onebit_compress(x_cpu, err_cpu)
    assert torch.equal(x.cpu(), x_cpu)  # input untouched
    assert torch.allclose(scale.cpu(), scale_ref, rtol=1e-5)
    assert (packed.cpu() != packed_ref).float().mean() < 1e-5  # ties at exactly 0 aside
    assert torch.allclose(err.cpu(), err_cpu, atol=1e-4)
    rows = 4
    pk = torch.randint(0, 256, (rows, n // 8 // rows), dtype=torch.uint8, device=_dev())
    sc = torch.rand(rows, device=_dev()) + 0.5
    for mode in (0, 1):
        got = onebit_decompress(pk, sc, mode)
        want = onebit_decompress(pk.cpu(), sc.cpu(), mode)
        assert torch.allclose(got.cpu(), want, atol=1e-6)


@pytest.mark.parametrize("batch_first", [True, False])
@pytest.mark.parametrize("dtype,hidden", [(torch.bfloat16, 4096), (torch.float32, 100), (torch.float16, 33)])
def test_random_ltd_token_ops(batch_first, dtype, hidden):
    from deepspeedsyclsupport_b200.ops.random_ltd import gpt_sample_tokens, token_gather, token_scatter_, token_sort_
    torch.manual_seed(0)
    B, S, R, L = 3, 257, 96, 2
    idx, _ = gpt_sample_tokens(R, S, B, layers=L, device=_dev())
    assert idx.dtype == torch.int32 and idx.shape == (L, B, R)
    assert bool((idx[..., 1:] > idx[..., :-1]).all()) and int(idx.min()) >= 0 and int(idx.max()) < S
    raw = torch.stack([torch.randperm(S, device=_dev())[:R] for _ in range(L * B)]).to(torch.int32).view(L, B, R).contiguous()
    assert torch.equal(token_sort_(raw.clone(), S), raw.sort(dim=-1)[0])
    x = torch.randn((B, S, hidden) if batch_first else (S, B, hidden), device=_dev()).to(dtype)
    sel = idx[0]
    got = token_gather(x, sel, batch_first)
    want = token_gather(x.cpu(), sel.cpu(), batch_first)
    assert torch.equal(got.cpu(), want)
    part = torch.randn_like(got)
    full = x.clone()
    token_scatter_(full, part, sel, batch_first)
    want_full = token_scatter_(x.cpu().clone(), part.cpu(), sel.cpu(), batch_first)
    assert torch.equal(full.cpu(), want_full)


def test_random_ltd_layer_autograd_gpu():
    from deepspeedsyclsupport_b200.ops.random_ltd import GatherTokens, ScatterTokens, gpt_sample_tokens
    torch.manual_seed(0)
    B, S, R, H = 2, 64, 16, 128
    idx, _ = gpt_sample_tokens(R, S, B, layers=1, device=_dev())
    x = torch.randn(B, S, H, device=_dev(), requires_grad=True)
    w = torch.randn(H, H, device=_dev(), requires_grad=True)
    full, part = GatherTokens.apply(x, idx[0], True)
    y = ScatterTokens.apply(full, part @ w, idx[0], True)
    y.square().sum().backward()
    xr = x.detach().clone().requires_grad_()
    wr = w.detach().clone().requires_grad_()
    ix = idx[0].long().unsqueeze(-1).expand(B, R, H)
    yr = xr.scatter(1, ix, torch.gather(xr, 1, ix) @ wr)
    yr.square().sum().backward()
    assert torch.allclose(y, yr, atol=1e-5) and torch.allclose(x.grad, xr.grad, atol=1e-3) and torch.allclose(w.grad, wr.grad, atol=1e-2)


@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16, torch.float32])
def test_nhwc_bias_add_kernel(dtype):
    from deepspeedsyclsupport_b200.ops.spatial import nhwc_bias_add, nhwc_bias_add_add, nhwc_bias_add_bias_add
    torch.manual_seed(0)
    x = torch.randn(2, 320, 17, 13, device=_dev(), dtype=dtype).contiguous(memory_format=torch.channels_last)
    o = torch.randn_like(x).contiguous(memory_format=torch.channels_last)
    b, ob = torch.randn(320, device=_dev(), dtype=dtype), torch.randn(320, device=_dev(), dtype=dtype)
    tol = 1e-5 if dtype == torch.float32 else 2e-2
    ref = x.float() + b.float().view(1, -1, 1, 1)
    assert torch.allclose(nhwc_bias_add(x, b).float(), ref, atol=tol)
    assert torch.allclose(nhwc_bias_add_add(x, b, o).float(), ref + o.float(), atol=tol)
    got = nhwc_bias_add_bias_add(x, b, o, ob)
    assert got.is_contiguous(memory_format=torch.channels_last)
    assert torch.allclose(got.float(), ref + o.float() + ob.float().view(1, -1, 1, 1), atol=2 * tol)


def test_transformer_layer_bf16_gpu():
    """DeepSpeedTransformerLayer on the tcgen05 GEMM + fused norm / bias-gelu kernels vs. its own fp32 CPU math."""
    from deepspeedsyclsupport_b200.ops.transformer import DeepSpeedTransformerConfig, DeepSpeedTransformerLayer
    torch.manual_seed(0)
    cfg = DeepSpeedTransformerConfig(batch_size=4, hidden_size=256, intermediate_size=1024, heads=4, attn_dropout_ratio=0.0, hidden_dropout_ratio=0.0,
                                     num_hidden_layers=2, initializer_range=0.02, pre_layer_norm=True, layer_norm_eps=1e-5)
    ref_layer = DeepSpeedTransformerLayer(cfg)
    ref_layer.inter_b.data.normal_(std=0.05)
    gpu_layer = DeepSpeedTransformerLayer(cfg)
    gpu_layer.load_state_dict(ref_layer.state_dict())
    gpu_layer = gpu_layer.to(_dev()).bfloat16()
    x = torch.randn(4, 128, 256)
    xg = x.to(_dev()).bfloat16().requires_grad_()
    xr = x.bfloat16().float().requires_grad_()
    y = gpu_layer(xg)
    yr = ref_layer(xr)
    assert (y.float().cpu() - yr).abs().max() < 0.08
    y.float().square().mean().backward()
    yr.square().mean().backward()
    g, gr = gpu_layer.inter_w.grad.float().cpu(), ref_layer.inter_w.grad
    assert (g - gr).abs().max() <= 0.05 * gr.abs().max() + 1e-5
    gb, gbr = gpu_layer.inter_b.grad.float().cpu(), ref_layer.inter_b.grad
    assert (gb - gbr).abs().max() <= 0.05 * gbr.abs().max() + 1e-5


@pytest.mark.parametrize("shape", [(8192, 64, 128), (40, 6, 64), (7, 5, 96)])
def test_rope_vectorised_matches_scalar_reference(shape):
    """large position ids (up to S) and several head-group splits: the vector kernel against the fp32 torch formula"""
    from deepspeedsyclsupport_b200.ops import fused_ops as FO
    torch.manual_seed(0)
    T, H, d = shape
    S = 4096
    inv = FO.rope_inv_freq(d, device=_dev())
    x = torch.randn(T, (H + 2) * d, device=_dev(), dtype=torch.bfloat16)
    pos = torch.randint(0, S, (T, ), device=_dev(), dtype=torch.int32)
    ref = FO.rope_ref(x.float(), inv, pos, 1, H, d)
    y = x.clone()
    FO._rope_apply(y, inv, pos, S, 1, H, d, False)
    assert (y.float() - ref).abs().max() < 4e-2
    assert torch.equal(y[:, :d], x[:, :d]) and torch.equal(y[:, (H + 1) * d:], x[:, (H + 1) * d:])  # untouched heads
    FO._rope_apply(y, inv, pos, S, 1, H, d, True)  # inverse rotation restores the input (up to bf16 rounding)
    assert (y.float() - x.float()).abs().max() < 6e-2


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16, torch.float32])
def test_sumsq(dtype):
    from deepspeedsyclsupport_b200.ops.fused_ops import sumsq
    torch.manual_seed(0)
    x = torch.randn(3_000_011, device=_dev()).to(dtype)
    got = sumsq(x)
    ref = x.double().square().sum()
    assert abs(got.item() - ref.item()) <= 1e-4 * ref.item()
    got2 = sumsq(x[:8 * 1000], out=got)
    assert abs(got2.item() - (ref + x[:8000].double().square().sum()).item()) <= 1e-4 * ref.item()


def test_checkpoint_policies_skip_kernels_and_keep_gradients():
    """Selective recomputation (ops/recompute.py): "attention" runs the flash forward once per layer and drops the unused
    down-projection recompute; "gemm" also replays the kept qkv / o_proj / gate_up outputs.  Gradients agree with full
    recomputation to within the backward's own run-to-run noise (its dQ reduction order is not deterministic)."""
    from deepspeedsyclsupport_b200.models import LlamaForCausalLM, llama_config
    from deepspeedsyclsupport_b200.ops import flash_attention as FA
    from deepspeedsyclsupport_b200.ops import gemm as G
    ids = torch.randint(0, 512, (2, 256), device=_dev())

    def run(policy):
        torch.manual_seed(0)
        cfg = llama_config("tiny", activation_checkpointing=True, checkpoint_policy=policy, max_position_embeddings=256)
        model = LlamaForCausalLM(cfg).to(_dev()).bfloat16().train()
        n = {"attn": 0, "gemm": 0}
        orig_attn, orig_gemm = FA.flash_fwd, G.gemm

        def count_attn(*a, **k):
            n["attn"] += 1
            return orig_attn(*a, **k)

        def count_gemm(*a, **k):
            n["gemm"] += 1
            return orig_gemm(*a, **k)

        FA.flash_fwd, G.gemm = count_attn, count_gemm
        try:
            model(ids, ids).backward()
        finally:
            FA.flash_fwd, G.gemm = orig_attn, orig_gemm
        return n, [p.grad.float().clone() for p in model.parameters()], cfg.num_hidden_layers

    n_full, g_full, L = run("full")
    _, g_full2, _ = run("full")
    n_attn, g_attn, _ = run("attention")
    n_gemm, g_gemm, _ = run("gemm")
    n_none, g_none, _ = run("none")
    assert n_full["attn"] == 2 * L and n_attn["attn"] == L and n_gemm["attn"] == L
    assert n_none["attn"] == L and n_none["gemm"] == n_full["gemm"] - 4 * L  # "none": nothing is recomputed
    for a, a2, d in zip(g_full, g_full2, g_none):
        assert (a - d).abs().max() <= 4 * (a - a2).abs().max() + 1e-3 * a.abs().max() + 1e-7
    assert n_attn["gemm"] == n_full["gemm"] - L          # down projection not recomputed
    assert n_gemm["gemm"] == n_full["gemm"] - 4 * L      # no GEMM at all during recomputation
    for a, a2, b, c in zip(g_full, g_full2, g_attn, g_gemm):
        noise = (a - a2).abs().max()
        tol = 4 * noise + 1e-3 * a.abs().max() + 1e-7
        assert (a - b).abs().max() <= tol and (a - c).abs().max() <= tol


def test_gemm_odd_output_width():
    """N not a multiple of 8 (a 97- or 50257-entry vocabulary): the wrapper pads the row stride, TMA clips the columns."""
    from deepspeedsyclsupport_b200.ops.gemm import gemm
    torch.manual_seed(0)
    a = torch.randn(150, 64, device=_dev(), dtype=torch.bfloat16)
    b = torch.randn(97, 64, device=_dev(), dtype=torch.bfloat16)
    c = gemm(a, b)
    ref = a.float() @ b.float().t()
    assert c.shape == (150, 97) and (c.float() - ref).abs().max() <= 1e-2 * ref.abs().max()
    c32 = gemm(a, b, out_dtype=torch.float32)
    assert (c32 - ref).abs().max() <= 1e-4 * ref.abs().max()


@pytest.mark.parametrize("case", [
    # (rows per expert, N, K)
    ([5, 0, 131, 128, 1, 300, 0, 77], 384, 256),      # empties, exact tile, tails, N = 3 n-blocks of 128
    ([1, 1, 1, 1], 264, 64),                          # decode-sized, N not a multiple of the n-block
    ([700, 513, 1024, 260], 1024, 512),               # wide-tile variant (>= 256 rows per expert on average)
    ([0, 0, 0, 9], 128, 128),
])
def test_moe_grouped_gemm(case):
    from deepspeedsyclsupport_b200.ops.gemm import moe_gemm, moe_gemm_ref
    rows, N, K = case
    E, R = len(rows), sum(rows)
    torch.manual_seed(0)
    pad = 37  # rows past cumsum[-1] (capacity padding) must stay untouched
    x = torch.randn(R + pad, K, device=_dev(), dtype=torch.bfloat16)
    w = torch.randn(E, N, K, device=_dev(), dtype=torch.bfloat16) * 0.1
    cum = torch.tensor(rows, device=_dev()).cumsum(0)
    out = torch.full((R + pad, N), 7.0, device=_dev(), dtype=torch.bfloat16)
    moe_gemm(x, w, cum, out=out)
    ref = moe_gemm_ref(x.float(), w.float(), cum)
    assert (out[:R].float() - ref[:R]).abs().max() <= 1e-2 * ref.abs().max()
    assert torch.all(out[R:] == 7.0)


def test_ragged_mixtral_engine_bf16_gpu():
    """v2 engine, Mixtral family on the device: gating / scatter / grouped GEMM / gather kernels vs the fp32 dense mixture."""
    from test_inference_v2 import MOE_CFG, _dense_mixtral_logits, _hf_mixtral_state
    from deepspeedsyclsupport_b200.inference.v2 import RaggedInferenceEngineConfig, build_engine_from_state_dict
    cfg = dict(MOE_CFG, hidden_size=256, intermediate_size=512, num_attention_heads=4, num_key_value_heads=2)
    ecfg = RaggedInferenceEngineConfig(state_manager={"max_ragged_batch_size": 256, "max_ragged_sequence_count": 8, "max_context": 256,
                                                      "memory_config": {"mode": "allocate", "size": 40}})
    sd = _hf_mixtral_state(cfg)
    eng = build_engine_from_state_dict("mixtral", cfg, sd, ecfg, dtype=torch.bfloat16)
    torch.manual_seed(4)
    a, b = torch.randint(0, 97, (150, )), torch.randint(0, 97, (6, ))
    ref_a, ref_b = _dense_mixtral_logits(sd, cfg, a), _dense_mixtral_logits(sd, cfg, b)
    out = eng.put([1, 2], [a[:149], b])
    # the same engine with the experts evaluated by a per-expert loop: isolates the grouped kernel from bf16 routing flips
    import deepspeedsyclsupport_b200.inference.v2.modules.implementations.moe.cutlass_multi_gemm as MOE
    from deepspeedsyclsupport_b200.ops.gemm import moe_gemm_ref
    assert type(eng._model.moe_mod).__name__ == "DSMultiGemmMoE"
    grouped = MOE.moe_gemm
    MOE.moe_gemm = lambda x, w, c, out=None: moe_gemm_ref(x, w, c)
    try:
        eng2 = build_engine_from_state_dict("mixtral", cfg, sd, ecfg, dtype=torch.bfloat16)
        out2 = eng2.put([1, 2], [a[:149], b])
    finally:
        MOE.moe_gemm = grouped
    scale = ref_a.abs().max()
    assert (out[0].float() - out2[0].float()).abs().max() < 0.02 * scale and (out[1].float() - out2[1].float()).abs().max() < 0.02 * scale
    # against the fp32 dense mixture (a near-tie in the bf16 router may pick another expert for sequence b's last token)
    assert (out[0].float().cpu() - ref_a[148]).abs().max() < 0.06 * scale
    assert (out[1].float().cpu() - ref_b[5]).abs().max() < 0.15 * scale
    out = eng.put([1], [a[149:150]])
    assert (out[0].float().cpu() - ref_a[149]).abs().max() < 0.06 * scale


def test_cuda_graph_wrappers_replay_matches_eager():
    """DSUNet / DSVAE (model_implementations/): the captured graph is replayed with refilled static inputs, one graph per input
    signature, and reproduces the eager forward."""
    from test_misc_components import _FakeUNet, _FakeVAE
    from deepspeedsyclsupport_b200.model_implementations import DSVAE, DSUNet
    torch.manual_seed(0)
    unet, vae = _FakeUNet().to(_dev()), _FakeVAE().to(_dev())
    w = DSUNet(unet, enable_cuda_graph=True)
    for step in range(4):
        x, t, e = torch.randn(2, 4, 8, 8, device=_dev()), torch.rand(2, device=_dev()), torch.randn(2, 5, 8, device=_dev())
        got = w(x, t, e).clone()
        assert torch.allclose(got, unet(x.contiguous(memory_format=torch.channels_last), t, e), atol=1e-5)
    assert w.cuda_graph_created and len(w._graphs) == 1
    w(torch.randn(1, 4, 8, 8, device=_dev()), torch.rand(1, device=_dev()), torch.randn(1, 5, 8, device=_dev()))
    assert len(w._graphs) == 2  # a new batch size gets its own graph
    v = DSVAE(vae)
    z = torch.randn(2, 4, 8, 8, device=_dev())
    assert torch.allclose(v.decode(z), vae.decode(z), atol=1e-5) and torch.allclose(v.decode(z * 2), vae.decode(z * 2), atol=1e-5)


def test_quantized_linear_on_device_matches_host_format():
    """inference/quantization on the device: the kernels' int8 / int4 payload equals the host restatement bit for bit, and the
    bf16 forward (dequantise -> tcgen05 GEMM) stays within quantization error of the dense layer."""
    from deepspeedsyclsupport_b200.inference.quantization import QuantizedLinear
    from deepspeedsyclsupport_b200.inference.quantization.utils import Quantizer
    torch.manual_seed(0)
    lin = torch.nn.Linear(256, 384).to(_dev()).bfloat16()
    x = torch.randn(50, 256, device=_dev(), dtype=torch.bfloat16)
    ref = lin(x).float()
    for bits, tol in ((8, 0.03), (4, 0.3)):
        cfg = {"num_bits": bits, "group_size": 64, "group_dim": 1, "symmetric": False}
        pd, sd, _ = Quantizer(cfg).quantize(lin.weight.data)
        ph, sh, _ = Quantizer(cfg).quantize(lin.weight.data.cpu())
        assert torch.allclose(sd.cpu(), sh, rtol=1e-5, atol=1e-8)
        # same storage format: the host restatement decodes the device payload; values differ by at most one step where
        # (x - min) * (1 / scale) and (x - min) / scale round differently
        from deepspeedsyclsupport_b200.inference.quantization.utils import dequantize_host
        wd = dequantize_host(pd.cpu(), sd.cpu(), bits, False, torch.float32)
        wh = dequantize_host(ph, sh, bits, False, torch.float32)
        step = sh[:, 0].repeat_interleave(64)
        assert ((wd - wh).abs() <= step * 1.001).all() and ((wd - wh).abs() > 0.5 * step).float().mean() < 0.02
        q = QuantizedLinear(cfg, lin)
        assert q.qweight.numel() == 256 * 384 * bits // 8
        out = q(x).float()
        assert (out - ref).abs().max() < tol * ref.abs().max()


@pytest.mark.parametrize("shape", [(256, 512, 256), (1000, 1096, 336), (4096, 4096, 4096), (77, 264, 128)])
def test_gemm_fp8_rowwise_scaled(shape):
    """fp8 (e4m3) tcgen05 GEMM with per-row / per-column scales against the fp32 product of the SAME quantised operands (exact up
    to accumulation order and the bf16 output rounding), and against the unquantised product within fp8 tolerance."""
    from deepspeedsyclsupport_b200.ops.gemm import gemm_fp8, quantize_fp8_rowwise
    M, N, K = shape
    torch.manual_seed(0)
    a = torch.randn(M, K, device=_dev()) * torch.rand(M, 1, device=_dev()) * 3
    b = torch.randn(N, K, device=_dev()) * 0.05
    aq, sa = quantize_fp8_rowwise(a)
    bq, sb = quantize_fp8_rowwise(b)
    c = gemm_fp8(aq, bq, sa, sb).float()
    ref_q = (aq.float() @ bq.float().t()) * sa[:, None] * sb[None, :]
    assert (c - ref_q).abs().max() <= 1e-2 * ref_q.abs().max()
    ref = a @ b.t()
    assert (c - ref).abs().max() <= 0.08 * ref.abs().max()
    c1 = gemm_fp8(aq, bq).float()  # no scales
    assert (c1 - aq.float() @ bq.float().t()).abs().max() <= 1e-2 * (aq.float() @ bq.float().t()).abs().max()


def test_ragged_engine_w8a8_gpu():
    """FastGen with quantization_mode = wf8af8 on the device: fp8 tensor-core projections, logits close to the bf16 engine."""
    from test_inference_v2 import _hf_llama_state
    from deepspeedsyclsupport_b200.inference.v2 import RaggedInferenceEngineConfig, build_engine_from_state_dict
    cfg = dict(vocab_size=512, hidden_size=512, intermediate_size=1024, num_hidden_layers=2, num_attention_heads=4, num_key_value_heads=2,
               rms_norm_eps=1e-5)
    sd = _hf_llama_state(cfg)
    mk = lambda mode: RaggedInferenceEngineConfig(state_manager={"max_ragged_batch_size": 256, "max_ragged_sequence_count": 8, "max_context": 256,
                                                                 "memory_config": {"mode": "allocate", "size": 40}},
                                                  quantization={"quantization_mode": mode})
    torch.manual_seed(2)
    a = torch.randint(0, 512, (130, ))
    ref = build_engine_from_state_dict("llama", cfg, sd, mk(""), dtype=torch.bfloat16).put([1], [a])[0].float()
    eng = build_engine_from_state_dict("llama", cfg, sd, mk("wf8af8"), dtype=torch.bfloat16)
    assert eng._model.layers[0]["mlp1_w"].dtype == torch.float8_e4m3fn and eng._model.layers[0]["mlp1_w"].scales.is_cuda
    out = eng.put([1], [a])[0].float()
    # the same quantised model evaluated on the host in fp32 (restated fp8 product): isolates the kernels from fp8 noise
    from deepspeedsyclsupport_b200.inference.v2.checkpoint import InMemoryModelEngine
    from deepspeedsyclsupport_b200.inference.v2.engine_factory import build_model_from_checkpoint
    host = build_model_from_checkpoint("llama", cfg, InMemoryModelEngine(sd), mk("wf8af8"), device=torch.device("cpu"), dtype=torch.float32)
    w_dev, w_host = eng._model.layers[1]["mlp2_w"], host.layers[1]["mlp2_w"]
    assert (w_dev.float().cpu() - w_host.float()).abs().max() <= 0.13 * w_host.float().abs().max()  # bf16 vs fp32 source weights
    assert (out - ref).abs().max() < 0.35 * ref.abs().max()  # fp8 noise of four projections per layer on random weights
    assert (out - ref).abs().mean() < 0.08 * ref.abs().max()


@pytest.mark.parametrize("mode,cls", [("wf8af8_block", "FP8BlockLinear"), ("wf8af8_mx", "MXFP8Linear")])
def test_ragged_engine_block_scaled_fp8_gpu(mode, cls):
    """FastGen with the block-scaled fp8 modes: wf8af8_block (1 x 128 fp32 scales, promotion in registers) and wf8af8_mx (MXFP8 on
    the hardware block-scaled MMA); logits close to the bf16 engine."""
    from test_inference_v2 import _hf_llama_state
    from deepspeedsyclsupport_b200.inference.v2 import RaggedInferenceEngineConfig, build_engine_from_state_dict
    cfg = dict(vocab_size=512, hidden_size=512, intermediate_size=1024, num_hidden_layers=2, num_attention_heads=4, num_key_value_heads=2,
               rms_norm_eps=1e-5)
    sd = _hf_llama_state(cfg)
    mk = lambda mode: RaggedInferenceEngineConfig(state_manager={"max_ragged_batch_size": 256, "max_ragged_sequence_count": 8, "max_context": 256,
                                                                 "memory_config": {"mode": "allocate", "size": 40}},
                                                  quantization={"quantization_mode": mode})
    torch.manual_seed(2)
    a = torch.randint(0, 512, (130, ))
    ref = build_engine_from_state_dict("llama", cfg, sd, mk(""), dtype=torch.bfloat16).put([1], [a])[0].float()
    eng = build_engine_from_state_dict("llama", cfg, sd, mk(mode), dtype=torch.bfloat16)
    w = eng._model.layers[0]["mlp1_w"]
    assert type(eng._model.qkv_lin).__name__ == cls and w.dtype == torch.uint8 and w.scales.is_cuda
    out = eng.put([1], [a])[0].float()
    assert (out - ref).abs().max() < 0.35 * ref.abs().max()
    assert (out - ref).abs().mean() < 0.08 * ref.abs().max()


@pytest.mark.parametrize("shape", [(300, 512, 256), (128, 4096, 1024), (1000, 1096, 328)])
def test_fused_gemm_allreduce_single_rank(shape):
    """GEMM -> all-reduce kernel at world 1 (the reducers wait on this rank's own tile flags only): equals the plain product,
    with and without the residual, over repeated calls (epoch / parity handling)."""
    from deepspeedsyclsupport_b200.parallel.fused import FusedGemmAllReduce
    from deepspeedsyclsupport_b200.parallel.symm import EmulatedPeers
    M, N, K = shape
    em = EmulatedPeers(1, FusedGemmAllReduce.bytes_needed(M, N) + 4096)
    op = FusedGemmAllReduce(em, M, N)
    torch.manual_seed(0)
    w = (torch.randn(N, K, device=_dev()) * 0.05).bfloat16()
    for it in range(3):
        x = torch.randn(M, K, device=_dev()).bfloat16()
        res = torch.randn(M, N, device=_dev()).bfloat16() if it == 1 else None
        out = op(x, w, residual=res).float()
        ref = x.float() @ w.float().t()
        # the partial is rounded to bf16 before the reduction (as it is when it crosses NVLink)
        want = ref.bfloat16().float() + (res.float() if res is not None else 0)
        assert (out - want).abs().max() <= 1e-2 * want.abs().max()


def test_grouped_experts_training_path_gpu(monkeypatch):
    """MoE training experts on the grouped tcgen05 GEMM (forward and dgrad) against the per-expert loop, bf16."""
    from torch import nn
    from deepspeedsyclsupport_b200.moe.experts import Experts

    class MLP(nn.Module):

        def __init__(self):
            super().__init__()
            self.up, self.down = nn.Linear(256, 512), nn.Linear(512, 256)

        def forward(self, x):
            return self.down(torch.nn.functional.gelu(self.up(x)))

    torch.manual_seed(0)
    grouped = Experts(MLP(), num_local_experts=4).to(_dev(), torch.bfloat16)
    assert grouped._grouped_act is not None
    with torch.no_grad():
        for e in grouped.deepspeed_experts:
            for p in e.parameters():
                p.add_(0.05 * torch.randn_like(p))
    monkeypatch.setenv("DSB200_GROUPED_EXPERTS", "0")
    loop = Experts(MLP(), num_local_experts=4).to(_dev(), torch.bfloat16)
    loop.load_state_dict(grouped.state_dict())
    x = torch.randn(1, 4, 150, 256, device=_dev(), dtype=torch.bfloat16, requires_grad=True)  # 150 rows per expert: partial tiles
    x2 = x.detach().clone().requires_grad_(True)
    w = torch.randn_like(x)
    yg, yl = grouped(x), loop(x2)
    (yg * w).sum().backward()
    (yl * w).sum().backward()
    scale = yl.float().abs().max()
    assert (yg.float() - yl.float()).abs().max() <= 2e-2 * scale
    assert (x.grad.float() - x2.grad.float()).abs().max() <= 3e-2 * x2.grad.float().abs().max()
    for (n, p), (_, q) in zip(grouped.named_parameters(), loop.named_parameters()):
        assert (p.grad.float() - q.grad.float()).abs().max() <= 3e-2 * q.grad.float().abs().max() + 1e-3, n


def test_fused_gemm_reduce_scatter_rows_single_rank():
    """Reduce-scatter form of the fused GEMM kernel at world 1: only the requested row range is reduced and written (ranges that cut
    through tiles, start inside a tile, or cover everything)."""
    from deepspeedsyclsupport_b200.parallel.fused import FusedGemmAllReduce
    from deepspeedsyclsupport_b200.parallel.symm import EmulatedPeers
    M, N, K = 700, 512, 256
    em = EmulatedPeers(1, FusedGemmAllReduce.bytes_needed(M, N) + 4096)
    op = FusedGemmAllReduce(em, M, N)
    torch.manual_seed(0)
    x = torch.randn(M, K, device=_dev()).bfloat16()
    w = (torch.randn(N, K, device=_dev()) * 0.05).bfloat16()
    ref = (x.float() @ w.float().t()).bfloat16().float()
    for lo, hi in ((0, 350), (350, 700), (130, 131), (0, 700), (255, 513)):
        res = torch.randn(hi - lo, N, device=_dev()).bfloat16()
        out = torch.full((hi - lo + 1, N), 7.0, device=_dev(), dtype=torch.bfloat16)  # one guard row
        op(x, w, residual=res, out=out[:hi - lo], rows=(lo, hi))
        want = ref[lo:hi] + res.float()
        assert (out[:hi - lo].float() - want).abs().max() <= 1e-2 * want.abs().max(), (lo, hi)
        assert torch.all(out[hi - lo] == 7.0)


@pytest.mark.parametrize("shape,dtype", [((37, 4096), torch.bfloat16), ((5, 11008), torch.bfloat16), ((64, 1024), torch.float16), ((3, 264), torch.float32)])
def test_fp8_rowwise_quantize_kernel_matches_eager(shape, dtype):
    """One-kernel per-token fp8 quantisation (rows longer than the 2048-element register cache take the second read) against the
    eager abs / amax / div / clamp / cast chain it replaces."""
    from deepspeedsyclsupport_b200.ops.gemm import FP8_MAX, quantize_fp8_rowwise
    torch.manual_seed(1)
    x = (torch.randn(*shape, device=_dev()) * torch.rand(shape[0], 1, device=_dev()) * 4).to(dtype)
    x[0] = 0  # an all-zero row must not divide by zero
    q, s = quantize_fp8_rowwise(x)
    xf = x.float()
    s_ref = xf.abs().amax(-1).clamp_min(1e-12) / FP8_MAX
    q_ref = (xf / s_ref[:, None]).clamp(-FP8_MAX, FP8_MAX).to(torch.float8_e4m3fn)
    assert torch.allclose(s, s_ref, rtol=1e-6, atol=0)
    # identical up to the rounding of x * (1/scale) versus x / scale at e4m3 tie points
    mism = (q.float() != q_ref.float()).float().mean().item()
    assert mism < 2e-3, mism
    assert ((q.float() * s[:, None] - xf).abs() <= 0.07 * xf.abs().amax(-1, keepdim=True) + 1e-9).all()


@pytest.mark.parametrize("shape", [(256, 256, 256), (1000, 1096, 384), (4096, 4096, 4096), (77, 264, 128), (512, 11008, 4096)])
def test_gemm_fp8_block_scaled(shape):
    """Block-scaled fp8 tcgen05 GEMM (1 x 128 scale blocks on both operands, promotion in registers) against the fp32 product of
    the dequantised operands, and against the unquantised product within fp8 tolerance; operands come from the ZeRO++ block
    quantiser, i.e. a qwZ weight payload is consumed as it is."""
    from deepspeedsyclsupport_b200.ops.gemm import gemm_fp8_block, linear_fp8_block
    from deepspeedsyclsupport_b200.ops.quantizer import dequantize_fp8_block, quantize_fp8_block
    M, N, K = shape
    torch.manual_seed(0)
    a = (torch.randn(M, K, device=_dev()) * (1 + 3 * torch.rand(M, K // 128, device=_dev()).repeat_interleave(128, 1))).bfloat16()
    b = (torch.randn(N, K, device=_dev()) * 0.05 * (1 + 5 * torch.rand(N, K // 128, device=_dev()).repeat_interleave(128, 1))).bfloat16()
    aq, sa = quantize_fp8_block(a)
    bq, sb = quantize_fp8_block(b)
    c = gemm_fp8_block(aq.view(M, K), bq.view(N, K), sa, sb).float()
    a_dq = dequantize_fp8_block(aq, sa, dtype=torch.float32).view(M, K)
    b_dq = dequantize_fp8_block(bq, sb, dtype=torch.float32).view(N, K)
    ref_q = a_dq @ b_dq.t()
    assert (c - ref_q).abs().max() <= 1e-2 * ref_q.abs().max(), (c - ref_q).abs().max() / ref_q.abs().max()
    ref = a.float() @ b.float().t()
    assert (c - ref).norm() <= 0.06 * ref.norm()
    y = linear_fp8_block(a, bq.view(N, K), sb).float()
    assert (y - ref).norm() <= 0.06 * ref.norm()


@pytest.mark.parametrize("shape", [(128, 256, 128), (256, 512, 512), (1000, 1096, 384), (4096, 4096, 4096), (77, 264, 128)])
def test_gemm_mxfp8_hardware_block_scaled(shape):
    """MXFP8 GEMM on tcgen05.mma.kind::mxf8f6f4.block_scale (UE8M0 scale per 32 elements, staged smem -> TMEM with tcgen05.cp):
    quantiser against its restatement, GEMM against the fp32 product of the dequantised operands (exact up to accumulation
    order + bf16 output) with scales that vary strongly per 32-element block, and against the unquantised product."""
    from deepspeedsyclsupport_b200.ops.gemm import gemm_mxfp8, linear_mxfp8, mxfp8_unpack_scales, quantize_mxfp8, _mxfp8_quantize_ref
    M, N, K = shape
    torch.manual_seed(0)
    blk = lambda r: torch.exp2(torch.randint(-6, 7, (r, K // 32), device=_dev()).float()).repeat_interleave(32, 1)
    a = (torch.randn(M, K, device=_dev()) * blk(M)).bfloat16()
    b = (torch.randn(N, K, device=_dev()) * 0.05 * blk(N)).bfloat16()
    aq, sfa = quantize_mxfp8(a)
    bq, sfb = quantize_mxfp8(b)
    aq_ref, sfa_ref = _mxfp8_quantize_ref(a.cpu())
    assert torch.equal(mxfp8_unpack_scales(sfa, M, K).cpu(), mxfp8_unpack_scales(sfa_ref, M, K))
    assert (aq.cpu().view(torch.float8_e4m3fn).float() != aq_ref.view(torch.float8_e4m3fn).float()).float().mean() < 1e-3
    a_dq = aq.view(torch.float8_e4m3fn).float() * mxfp8_unpack_scales(sfa, M, K).repeat_interleave(32, 1)
    b_dq = bq.view(torch.float8_e4m3fn).float() * mxfp8_unpack_scales(sfb, N, K).repeat_interleave(32, 1)
    ref_q = a_dq @ b_dq.t()
    c = gemm_mxfp8(aq, bq, sfa, sfb).float()
    assert (c - ref_q).abs().max() <= 1e-2 * ref_q.abs().max(), ((c - ref_q).abs().max() / ref_q.abs().max()).item()
    ref = a.float() @ b.float().t()
    assert (c - ref).norm() <= 0.06 * ref.norm()
    y = linear_mxfp8(a, bq, sfb).float()
    assert (y - ref).norm() <= 0.06 * ref.norm()


@pytest.mark.parametrize("rotate_half,kv_heads,head_dim", [(True, 2, 64), (False, 8, 64), (True, 4, 128)])
def test_v1_attention_kv_arena_matches_dense_path(monkeypatch, rotate_half, kv_heads, head_dim):
    """Kernel-injection attention: prompt + decode steps on the KV arena (rotary/append kernel + tcgen05 paged flash) equal the
    dense masked-softmax path of the same module, including growth of the arena past its initial size."""
    from deepspeedsyclsupport_b200.ops.transformer.inference import DeepSpeedInferenceConfig
    from deepspeedsyclsupport_b200.ops.transformer.inference.ds_attention import DeepSpeedSelfAttention
    torch.manual_seed(0)
    heads = 8
    H = heads * head_dim
    cfg = DeepSpeedInferenceConfig(hidden_size=H, heads=heads, num_kv=kv_heads if kv_heads != heads else -1, dtype=torch.bfloat16,
                                   rotary_dim=head_dim // 2, rotate_half=rotate_half, max_out_tokens=128, triangular_masking=True)
    a, b = DeepSpeedSelfAttention(cfg), DeepSpeedSelfAttention(cfg)
    for name in ("attn_qkvw", "attn_qkvb", "attn_ow", "attn_ob"):
        w = torch.randn_like(getattr(a, name).float()) * (0.05 if name.endswith("w") else 0.1)
        getattr(a, name).data.copy_(w)
        getattr(b, name).data.copy_(w)
    steps = [torch.randn(2, 250, H, device="cuda").bfloat16()] + [torch.randn(2, 1, H, device="cuda").bfloat16() for _ in range(8)]
    outs = {}
    for tag, mod, flag in (("arena", a, "1"), ("dense", b, "0")):
        monkeypatch.setenv("DSB200_V1_ARENA", flag)
        outs[tag] = [mod(x.clone(), get_present=True) for x in steps]
    assert a.arena.length == 258 and a.arena.per_seq == 4 and a.k_cache is None and b.arena.length == 0  # grew from 2 blocks
    for (oa, ka, va), (ob, kb, vb) in zip(outs["arena"], outs["dense"]):
        assert oa.shape == ob.shape and ka.shape == kb.shape
        assert (oa.float() - ob.float()).abs().max() < 0.03 * ob.float().abs().max() + 1e-3
        assert (ka.float() - kb.float()).abs().max() < 0.07 and (va.float() - vb.float()).abs().max() < 1e-6  # rotated k: a bf16 ulp or two


@pytest.mark.parametrize("M,N,K", [(1, 5120, 8192), (16, 8192, 3584), (33, 1024, 4096), (64, 1280, 8192), (128, 2048, 4096), (7, 1000, 520), (32, 16384, 512)])
def test_gemm_skinny_weight_streaming(M, N, K):
    """Decode-sized GEMM (csrc/gemm/gemm_skinny.cu: swapped roles, split K, fp32 partials + reduce) vs an fp32 matmul of the same
    bf16 operands; `gemm()` must route these shapes to it."""
    from deepspeedsyclsupport_b200.ops import gemm as G
    from deepspeedsyclsupport_b200.ops import native
    torch.manual_seed(M + N)
    x = (torch.randn(M, K, device="cuda") * 0.5).bfloat16()
    w = (torch.randn(N, K, device="cuda") * 0.05).bfloat16()
    before = native.launch_count
    y = G.gemm(x, w)
    assert native.launch_count > before
    ref = x.float() @ w.float().t()
    err = (y.float() - ref).abs().max().item()
    assert y.shape == (M, N) and err < 0.01 * ref.abs().max().item() + 1e-2, (M, N, K, err)
    # strided output view (a column slice of a wider buffer), as the fused-QKV callers use
    wide = torch.zeros(M, N + 64, device="cuda", dtype=torch.bfloat16)
    G.gemm(x, w, out=wide[:, :N])
    assert (wide[:, :N].float() - ref).abs().max().item() < 0.01 * ref.abs().max().item() + 1e-2 and float(wide[:, N:].abs().max()) == 0.0


@pytest.mark.parametrize("bits,M,N,K,group", [(8, 16, 4096, 4096, 128), (4, 16, 4096, 4096, 64), (8, 1, 1024, 8192, 8192), (4, 100, 2048, 1024, 256),
                                             (8, 128, 11008, 512, 64)])
def test_mixed_gemm_weight_only_quantised_skinny(bits, M, N, K, group):
    """MixedGEMM at decode size runs the mixed-input weight-streaming kernel (int8 / packed int4 weights dequantised in shared
    memory, tcgen05 MMA); oracle = fp32 matmul against the group-wise dequantised weight."""
    from deepspeedsyclsupport_b200.inference.v2.inference_utils import ActivationType
    from deepspeedsyclsupport_b200.inference.v2.kernels.cutlass_ops import MixedGEMM, _dequant, _mixed_skinny
    torch.manual_seed(bits * 1000 + M)
    x = (torch.randn(M, K, device="cuda") * 0.5).bfloat16()
    sc = (torch.rand(N, K // group, device="cuda") * 0.01 + 0.002)
    if bits == 8:
        wq = torch.randint(-127, 128, (N, K), device="cuda", dtype=torch.int8)
    else:
        q4 = torch.randint(0, 16, (N, K), device="cuda")
        wq = (q4[:, 0::2] | (q4[:, 1::2] << 4)).to(torch.uint8)
    assert _mixed_skinny(x, wq, sc, bits) is not None, "the fused kernel did not take this shape"
    out = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
    MixedGEMM("bf16", ActivationType.IDENTITY, bits)(out, x, wq, sc)
    ref = x.float() @ _dequant(wq, sc, bits, torch.float32).t()
    err = (out.float() - ref).abs().max().item()
    assert err < 0.01 * ref.abs().max().item() + 1e-2, (bits, M, N, K, group, err)


@pytest.mark.parametrize("B,N,L,H,D,dtype,use_b1,use_b2", [(1, 3, 70, 2, 32, torch.bfloat16, True, True), (2, 2, 33, 3, 64, torch.float16, True, False),
                                                         (1, 2, 128, 4, 16, torch.bfloat16, False, True), (1, 1, 17, 1, 8, torch.bfloat16, False, False)])
def test_evoformer_attention_kernels(B, N, L, H, D, dtype, use_b1, use_b2):
    """csrc/attn/evoformer.cu (fused forward + LSE, row / column backward passes with bias gradients) against the explicit fp32
    formula and its autograd gradients; bias1 carries -inf style mask entries."""
    import math
    from deepspeedsyclsupport_b200.ops.deepspeed4science.evoformer_attn import DS4Sci_EvoformerAttention, _native_ok
    torch.manual_seed(L + D)
    dev = "cuda"
    q, k, v = (torch.randn(B, N, L, H, D, device=dev).to(dtype).requires_grad_(True) for _ in range(3))
    biases = []
    b1 = b2 = None
    if use_b1:
        mask = torch.rand(B, N, 1, 1, L, device=dev) < 0.2
        mask[..., 0] = False  # never a fully masked row
        b1 = (torch.randn(B, N, 1, 1, L, device=dev) * 0.5).masked_fill(mask, -1e4).to(dtype).requires_grad_(True)
    if use_b2:
        b2 = torch.randn(B, 1, H, L, L, device=dev).to(dtype).requires_grad_(True)
    biases = [b for b in (b1, b2) if b is not None] if (use_b1 or not use_b2) else [None, b2]
    assert _native_ok(q, k, v, b1, b2)
    out = DS4Sci_EvoformerAttention(q, k, v, biases)
    g = torch.randn_like(out)
    out.backward(g)
    # fp32 oracle on the same (rounded) inputs
    qf, kf, vf = (t.detach().float().requires_grad_(True) for t in (q, k, v))
    b1f = None if b1 is None else b1.detach().float().requires_grad_(True)
    b2f = None if b2 is None else b2.detach().float().requires_grad_(True)
    s = torch.einsum("bnihd,bnjhd->bnhij", qf, kf) / math.sqrt(D)
    if b1f is not None:
        s = s + b1f  # [B, N, 1, 1, L] broadcasts over heads and query rows
    if b2f is not None:
        s = s + b2f  # [B, 1, H, L, L] broadcasts over N
    ref = torch.einsum("bnhij,bnjhd->bnihd", torch.softmax(s, -1), vf)
    ref.backward(g.float())
    tol = 2e-2 if dtype == torch.bfloat16 else 4e-3

    def close(a, b_, name):
        err = (a.float() - b_).abs().max().item()
        assert err <= tol * max(1.0, b_.abs().max().item()), (name, err, b_.abs().max().item())

    close(out, ref, "out")
    close(q.grad, qf.grad, "dq")
    close(k.grad, kf.grad, "dk")
    close(v.grad, vf.grad, "dv")
    if b1 is not None:
        close(b1.grad, b1f.grad, "dbias1")
    if b2 is not None:
        close(b2.grad, b2f.grad, "dbias2")


def test_v1_attention_kv_arena_with_alibi(monkeypatch):
    """BLOOM-style ALiBi on the arena path: the per-head slopes are read off the alibi tensor and applied inside the paged flash
    kernels; prompt + decode steps equal the dense masked-softmax path that adds the tensor itself."""
    from deepspeedsyclsupport_b200.ops.transformer.inference import DeepSpeedInferenceConfig
    from deepspeedsyclsupport_b200.ops.transformer.inference.ds_attention import DeepSpeedSelfAttention
    torch.manual_seed(3)
    heads, head_dim = 8, 64
    H = heads * head_dim
    cfg = DeepSpeedInferenceConfig(hidden_size=H, heads=heads, dtype=torch.bfloat16, max_out_tokens=256, triangular_masking=True)
    a, b = DeepSpeedSelfAttention(cfg), DeepSpeedSelfAttention(cfg)
    for name in ("attn_qkvw", "attn_qkvb", "attn_ow", "attn_ob"):
        w = torch.randn_like(getattr(a, name).float()) * (0.05 if name.endswith("w") else 0.1)
        getattr(a, name).data.copy_(w)
        getattr(b, name).data.copy_(w)
    slopes = 2.0 ** (-8.0 * torch.arange(1, heads + 1, device="cuda") / heads)
    B = 1
    steps = [torch.randn(B, 90, H, device="cuda").bfloat16()] + [torch.randn(B, 1, H, device="cuda").bfloat16() for _ in range(4)]
    outs = {}
    for tag, mod, flag in (("arena", a, "1"), ("dense", b, "0")):
        monkeypatch.setenv("DSB200_V1_ARENA", flag)
        T, res = 0, []
        for x in steps:
            T += x.shape[1]
            alibi = (slopes[:, None] * torch.arange(T, device="cuda")[None, :]).bfloat16().view(B * heads, 1, T)  # as HF builds it
            res.append(mod(x.clone(), alibi=alibi)[0])
        outs[tag] = res
    assert a.arena.length == 94 and a._alibi_slopes is not None and b.arena.length == 0
    for oa, ob in zip(outs["arena"], outs["dense"]):
        assert (oa.float() - ob.float()).abs().max() < 0.04 * ob.float().abs().max() + 2e-3


def test_v1_attention_kv_arena_with_local_window(monkeypatch):
    """GPT-Neo style local attention (window 64) on the arena path equals the dense masked-softmax path, prompt longer than the
    window plus decode steps."""
    from deepspeedsyclsupport_b200.ops.transformer.inference import DeepSpeedInferenceConfig
    from deepspeedsyclsupport_b200.ops.transformer.inference.ds_attention import DeepSpeedSelfAttention
    torch.manual_seed(4)
    heads, head_dim = 4, 128
    H = heads * head_dim
    cfg = DeepSpeedInferenceConfig(hidden_size=H, heads=heads, dtype=torch.bfloat16, max_out_tokens=256, triangular_masking=True,
                                   local_attention=True, window_size=64)
    a, b = DeepSpeedSelfAttention(cfg), DeepSpeedSelfAttention(cfg)
    for name in ("attn_qkvw", "attn_qkvb", "attn_ow", "attn_ob"):
        w = torch.randn_like(getattr(a, name).float()) * (0.05 if name.endswith("w") else 0.1)
        getattr(a, name).data.copy_(w)
        getattr(b, name).data.copy_(w)
    steps = [torch.randn(2, 150, H, device="cuda").bfloat16()] + [torch.randn(2, 1, H, device="cuda").bfloat16() for _ in range(3)]
    outs = {}
    for tag, mod, flag in (("arena", a, "1"), ("dense", b, "0")):
        monkeypatch.setenv("DSB200_V1_ARENA", flag)
        outs[tag] = [mod(x.clone())[0] for x in steps]
    assert a.arena.length == 153 and b.arena.length == 0
    for oa, ob in zip(outs["arena"], outs["dense"]):
        assert (oa.float() - ob.float()).abs().max() < 0.04 * ob.float().abs().max() + 2e-3
