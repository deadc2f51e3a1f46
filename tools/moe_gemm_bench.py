"""Grouped expert GEMM (dsb200_moe_gemm_bf16) vs a per-expert loop of dense GEMMs with a host sync on the row counts
(what the path did before) on Mixtral-8x7B expert shapes.  CUDA events, L2 flushed between iterations."""
import json
import sys

import torch

sys.path.insert(0, ".")
from deepspeedsyclsupport_b200.ops.gemm import gemm, moe_gemm  # noqa: E402


def timeit(fn, flush, iters=10, warm=3):
    for _ in range(warm):
        fn()
    ts = []
    for _ in range(iters):
        flush.fill_(1.0)
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        s.record()
        fn()
        e.record()
        torch.cuda.synchronize()
        ts.append(s.elapsed_time(e))
    ts.sort()
    return ts[len(ts) // 2]


def main():
    dev = torch.device("cuda")
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev).float()[:64 << 20]
    E, K, N = 8, 4096, 28672  # gate_up of Mixtral-8x7B: [2 * 14336, 4096] per expert
    w = torch.randn(E, N, K, device=dev, dtype=torch.bfloat16) * 0.02
    out = []
    for tokens in (16, 256, 4096, 16384):
        R = tokens * 2  # top-2
        g = torch.Generator(device="cpu").manual_seed(tokens)
        assign = torch.randint(0, E, (R, ), generator=g)
        counts = torch.bincount(assign, minlength=E)
        cum = counts.cumsum(0).to(dev)
        x = torch.randn(R, K, device=dev, dtype=torch.bfloat16)
        y = torch.empty(R, N, device=dev, dtype=torch.bfloat16)

        def grouped():
            moe_gemm(x, w, cum, out=y)

        def loop():
            b = [0] + cum.tolist()
            for e in range(E):
                if b[e + 1] > b[e]:
                    gemm(x[b[e]:b[e + 1]], w[e], out=y[b[e]:b[e + 1]])

        tg, tl = timeit(grouped, flush), timeit(loop, flush)
        flops = 2.0 * R * N * K
        wbytes = E * N * K * 2
        out.append({"tokens": tokens, "rows": R, "grouped_ms": round(tg, 4), "loop_ms": round(tl, 4), "speedup": round(tl / tg, 2),
                    "grouped_tflops": round(flops / tg / 1e9, 1), "weight_stream_gbs": round(wbytes / tg / 1e6, 1)})
        print(out[-1], flush=True)
    print(json.dumps({"bench": "moe_gemm Mixtral-8x7B gate_up (E=8, N=28672, K=4096)", "results": out}))


if __name__ == "__main__":
    main()
