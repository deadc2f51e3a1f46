"""GPU self-check + timing of the flash attention kernels against torch SDPA."""
import math
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from deepspeedsyclsupport_b200.ops.attention import attention_ref
from deepspeedsyclsupport_b200.ops import flash_attention as FA


def timeit(fn, iters=10):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(iters):
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        fn()
        e.record()
        torch.cuda.synchronize()
        ts.append(s.elapsed_time(e))
    return sorted(ts)[len(ts) // 2]


def main():
    bwd = "--bwd" in sys.argv
    torch.manual_seed(0)
    for (B, S, hq, hkv, d, causal) in [(1, 128, 1, 1, 128, True), (2, 256, 2, 2, 128, True), (2, 512, 4, 2, 128, True),
                                       (1, 384, 2, 2, 64, True), (2, 256, 2, 1, 128, False), (1, 2048, 4, 4, 128, True)]:
        qkv = torch.randn(B * S, (hq + 2 * hkv) * d, device="cuda", dtype=torch.bfloat16)
        out, lse = FA.flash_fwd(qkv, B, S, hq, hkv, d, causal)
        torch.cuda.synchronize()
        qf = qkv.float().requires_grad_()
        ref = attention_ref(qf, B, S, hq, hkv, d, causal)
        err = (out.float() - ref).abs().max().item()
        q, k, _ = qf.detach().view(B, S, hq + 2 * hkv, d).split([hq, hkv, hkv], dim=2)
        sc = torch.einsum("bshd,bthd->bhst", q, k.repeat_interleave(hq // hkv, 2)) / math.sqrt(d)
        if causal:
            sc = sc.masked_fill(torch.triu(torch.ones(S, S, device="cuda", dtype=torch.bool), 1), float("-inf"))
        lse_ref = torch.logsumexp(sc, -1)
        lerr = (lse - lse_ref).abs().max().item()
        print(f"FWD B={B} S={S} hq={hq} hkv={hkv} d={d} causal={causal}: out max_err={err:.4f} lse max_err={lerr:.5f}", flush=True)
        if bwd:
            dout = torch.randn_like(out)
            dqkv = FA.flash_bwd(dout, qkv, out, lse, B, S, hq, hkv, d, causal)
            torch.cuda.synchronize()
            ref.backward(dout.float())
            g = qf.grad
            nq, nk = hq * d, hkv * d
            for name, a, b in (("dq", dqkv[:, :nq], g[:, :nq]), ("dk", dqkv[:, nq:nq + nk], g[:, nq:nq + nk]),
                               ("dv", dqkv[:, nq + nk:], g[:, nq + nk:])):
                e = (a.float() - b).abs().max().item()
                print(f"   BWD {name}: max_err={e:.4f} (ref max {b.abs().max().item():.3f})", flush=True)
    if "--bench" in sys.argv:
        for (B, S, hq, hkv, d) in [(2, 4096, 32, 32, 128), (1, 16384, 32, 32, 128), (2, 4096, 64, 8, 128)]:
            qkv = torch.randn(B * S, (hq + 2 * hkv) * d, device="cuda", dtype=torch.bfloat16)
            fl = 4.0 * B * hq * S * S * d / 2
            t = timeit(lambda: FA.flash_fwd(qkv, B, S, hq, hkv, d, True))
            q, k, v = qkv.view(B, S, hq + 2 * hkv, d).split([hq, hkv, hkv], dim=2)
            q, k, v = (t_.transpose(1, 2).contiguous() for t_ in (q, k, v))
            t2 = timeit(lambda: torch.nn.functional.scaled_dot_product_attention(q, k, v, is_causal=True, enable_gqa=hq != hkv))
            print(f"BENCH FWD B={B} S={S} hq={hq} hkv={hkv}: own {t:.3f} ms = {fl / t / 1e9:.0f} TFLOP/s | torch SDPA {t2:.3f} ms = {fl / t2 / 1e9:.0f} TFLOP/s", flush=True)
            if bwd:
                out, lse = FA.flash_fwd(qkv, B, S, hq, hkv, d, True)
                dout = torch.randn_like(out)
                tb = timeit(lambda: FA.flash_bwd(dout, qkv, out, lse, B, S, hq, hkv, d, True))
                qr, kr, vr = (t_.detach().requires_grad_() for t_ in (q, k, v))
                o2 = torch.nn.functional.scaled_dot_product_attention(qr, kr, vr, is_causal=True, enable_gqa=hq != hkv)
                do2 = torch.randn_like(o2)
                tb2 = timeit(lambda: torch.autograd.grad(o2, (qr, kr, vr), do2, retain_graph=True))
                print(f"BENCH BWD: own {tb:.3f} ms = {2.5 * fl / tb / 1e9:.0f} TFLOP/s | torch SDPA {tb2:.3f} ms = {2.5 * fl / tb2 / 1e9:.0f} TFLOP/s", flush=True)


if __name__ == "__main__":
    main()
