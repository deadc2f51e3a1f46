"""The peer-memory protocols exercised CONCURRENTLY on a single GPU: W processes share device 0, map each other's buffers through
CUDA IPC (gloo carries the handles), and run the device barrier, the pull / copy-engine collectives, the fp8 reduce-scatter, the
fused reduce-scatter + AdamW and the MoE push / pull kernels against each other -- the flag / epoch protocols really wait on
another process's kernel here (the GPU time-slices the processes), which the single-process EmulatedPeers tests cannot show.
Oracle: the same reductions computed on the host from the gathered inputs."""
import pytest
import torch

from common import run_distributed

pytestmark = pytest.mark.gpu


def _gather_cpu(t):
    from deepspeedsyclsupport_b200 import comm as dist
    parts = [None] * dist.get_world_size()
    dist.all_gather_object(parts, t.detach().float().cpu())
    return parts


def _body():
    import deepspeedsyclsupport_b200 as ds  # noqa: F401
    from deepspeedsyclsupport_b200 import comm as dist
    from deepspeedsyclsupport_b200.moe.peer_dispatch import PeerMoEBuffers  # noqa: F401  (registers the kernel signatures)
    from deepspeedsyclsupport_b200.ops import quantizer as Q
    from deepspeedsyclsupport_b200.parallel.symm import SymmetricBuffer
    torch.cuda.set_device(0)
    dev = torch.device("cuda", 0)
    rank, W = dist.get_rank(), dist.get_world_size()
    n = 128 * 64
    sb = SymmetricBuffer(dist.get_world_group(), 8 * W * n * 4, backend="ipc")
    assert sb.backend == "ipc" and len(set(sb.ptrs)) == W
    torch.manual_seed(50 + rank)
    for it in range(3):  # several rounds: epochs advance, buffers are reused right after the barriers
        grad = torch.randn(W * n, device=dev).bfloat16()
        sb.view(0, (W * n, ), torch.bfloat16).copy_(grad)
        sb.barrier()
        everyone = _gather_cpu(grad)
        want = torch.stack(everyone).sum(0)[rank * n:(rank + 1) * n] / W
        # pull reduce-scatter with scale + fused sum of squares
        sq = torch.zeros(1, device=dev)
        dst = torch.zeros(n, device=dev)
        sb.reduce_scatter(dst, 0, torch.bfloat16, n, scale=1.0 / W, sumsq=sq)
        torch.cuda.synchronize()
        assert (dst.cpu() - want).abs().max() < 1e-5 and abs(sq.item() - want.square().sum().item()) < 1e-3 * sq.item() + 1e-6
        # all-gather: pull kernel and copy engines
        for fn in (sb.all_gather, sb.all_gather_ce):
            out = torch.empty(W * n, device=dev, dtype=torch.bfloat16)
            fn(out, 0, n * 2)
            torch.cuda.synchronize()
            assert torch.equal(out.float().cpu(), torch.cat([e[:n] for e in everyone]))
        # qgZ: quantise into the symmetric buffer, barrier, fp8 pull reduce-scatter
        q_off, s_off = W * n * 2, W * n * 3
        Q.quantize_fp8_block(grad, out=(sb.view(q_off, (W * n, ), torch.uint8), sb.view(s_off, (W * n // 128, ), torch.float32)))
        sb.barrier()
        d8 = torch.zeros(n, device=dev)
        sb.fp8_reduce_scatter(d8, q_off, s_off, n, scale=1.0 / W)
        torch.cuda.synchronize()
        assert ((d8.cpu() - want).norm() / want.norm()).item() < 6e-2
        # fused reduce-scatter + AdamW on a sub-range of the slice
        p = torch.randn(n, device=dev)
        m, v = torch.zeros(n, device=dev), torch.zeros(n, device=dev)
        lp, g_out = torch.zeros(n, device=dev, dtype=torch.bfloat16), torch.zeros(n, device=dev, dtype=torch.bfloat16)
        lo, cnt = 256, n - 512
        p0 = p.clone()
        sb.rs_adam_range(0, n, lo, p[lo:lo + cnt], m[lo:lo + cnt], v[lo:lo + cnt], lp[lo:lo + cnt], scale=1.0 / W, lr=1e-2, beta1=0.9, beta2=0.95,
                         eps=1e-8, weight_decay=0.1, step=1, g_out=g_out[lo:lo + cnt])
        torch.cuda.synchronize()
        ref = torch.nn.Parameter(p0[lo:lo + cnt].cpu().clone())
        opt = torch.optim.AdamW([ref], lr=1e-2, betas=(0.9, 0.95), eps=1e-8, weight_decay=0.1)
        ref.grad = want[lo:lo + cnt].clone()
        opt.step()
        assert (p[lo:lo + cnt].cpu() - ref.data).abs().max() < 1e-5 and torch.equal(p[:lo].cpu(), p0[:lo].cpu())
        assert (g_out[lo:lo + cnt].float().cpu() - want[lo:lo + cnt]).abs().max() < 2e-2
        sb.barrier()
    # MoE dispatch push / combine pull: 2 local experts per rank, capacity 8, top-2
    S, K, M, n_local, cap = 24, 2, 64, 2, 16
    E = W * n_local
    g = torch.Generator().manual_seed(7 + rank)
    x = torch.randn(S, M, generator=g).bfloat16().to(dev)
    te = torch.stack([torch.randperm(E, generator=g)[:K] for _ in range(S)]).int()
    te[::5, 1] = -1  # some dropped second choices
    ts = torch.zeros(S, K, dtype=torch.int32)
    fill = {}
    for s_ in range(S):
        for k_ in range(K):
            e = int(te[s_, k_])
            if e >= 0:
                ts[s_, k_] = fill.get(e, 0)
                fill[e] = fill.get(e, 0) + 1
    assert max(fill.values()) <= cap
    tw = torch.rand(S, K, generator=g)
    te_d, ts_d, tw_d = te.to(dev), ts.to(dev), tw.to(dev)
    rows = W * n_local * cap
    mb = SymmetricBuffer(dist.get_world_group(), 2 * rows * M * 2, backend="ipc")
    tab = __import__("ctypes").cast((__import__("ctypes").c_void_p * W)(*mb.ptrs), __import__("ctypes").c_void_p)
    from deepspeedsyclsupport_b200.ops import native
    A = mb.view(0, (W, n_local, cap, M), torch.bfloat16)
    A.zero_()
    mb.barrier()
    native.check(native.lib().dsb200_moe_dispatch_push(tab, 0, native.ptr(x), native.ptr(te_d), native.ptr(ts_d), None, S, K, M, rank, W, n_local, cap,
                                                       native.stream_ptr()))
    mb.barrier()
    torch.cuda.synchronize()
    xs, tes, tss = _gather_cpu(x), _gather_cpu(te), _gather_cpu(ts)
    want_recv = torch.zeros(W, n_local, cap, M)
    for src in range(W):
        for s_ in range(S):
            for k_ in range(K):
                e = int(tes[src][s_, k_])
                if e >= 0 and e // n_local == rank:
                    want_recv[src, e % n_local, int(tss[src][s_, k_])] = xs[src][s_]
    assert torch.equal(A.float().cpu(), want_recv)
    # every rank serves 2 * its received rows; tokens pull them back weighted
    B = mb.view(rows * M * 2, (W, n_local, cap, M), torch.bfloat16)
    B.copy_((A.float() * 2).bfloat16())
    mb.barrier()
    y = torch.empty(S, M, device=dev, dtype=torch.bfloat16)
    native.check(native.lib().dsb200_moe_combine_pull(tab, rows * M * 2, native.ptr(y), native.ptr(te_d), native.ptr(ts_d), native.ptr(tw_d), None, None, S, K, M,
                                                      rank, W, n_local, cap, native.stream_ptr()))
    mb.barrier()
    torch.cuda.synchronize()
    want_y = ((te >= 0).float() * tw).sum(1, keepdim=True) * 2 * x.float().cpu()
    assert (y.float().cpu() - want_y).abs().max() < 3e-2 * want_y.abs().max()
    torch.cuda.synchronize()
    mb.close()
    sb.close()


@pytest.mark.parametrize("world", [2, 4])
def test_peer_protocols_between_processes_on_one_gpu(world):
    run_distributed(_body, world_size=world, backend="gloo", timeout=600)
