"""fp8 GEMMs -- row-scaled (dsb200_gemm_fp8) and block-scaled 1 x 128 (dsb200_gemm_fp8_block) -- vs the bf16 CTA-pair GEMM and
torch._scaled_mm (cuBLASLt fp8, row-wise scales) on the training shapes.  CUDA events, L2 flushed between iterations, clock record.
    python tools/gemm_fp8_bench.py [--out profiles/gemm_fp8_bench_r2.json]"""
import json
import sys

import torch

sys.path.insert(0, ".")
from bench import ClockSampler  # noqa: E402
from deepspeedsyclsupport_b200.ops.gemm import gemm, gemm_fp8, gemm_fp8_block, gemm_mxfp8, quantize_fp8_rowwise, quantize_mxfp8  # noqa: E402
from deepspeedsyclsupport_b200.ops.quantizer import quantize_fp8_block  # noqa: E402


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
    flush = torch.empty(64 << 20, dtype=torch.float32, device=dev)
    sampler = ClockSampler(0)
    sampler.start()
    rows = []
    for M, N, K in ((8192, 4096, 4096), (8192, 22016, 4096), (8192, 4096, 11008), (8192, 8192, 8192)):
        a = torch.randn(M, K, device=dev)
        b = torch.randn(N, K, device=dev) * 0.05
        aq, sa = quantize_fp8_rowwise(a)
        bq, sb = quantize_fp8_rowwise(b)
        ab, bb = a.bfloat16(), b.bfloat16()
        out = torch.empty(M, N, device=dev, dtype=torch.bfloat16)
        t8 = timeit(lambda: gemm_fp8(aq, bq, sa, sb, out=out), flush)
        t16 = timeit(lambda: gemm(ab, bb, out=out), flush)
        try:
            tl = timeit(lambda: torch._scaled_mm(aq, bq.t(), scale_a=sa[:, None], scale_b=sb[None, :], out_dtype=torch.bfloat16), flush)
        except Exception as ex:  # noqa: BLE001
            tl = float("nan")
            print("torch._scaled_mm unavailable:", str(ex)[:80])
        tb = float("nan")
        if K % 128 == 0:
            aqb, sab = quantize_fp8_block(ab)
            bqb, sbb = quantize_fp8_block(bb)
            tb = timeit(lambda: gemm_fp8_block(aqb.view(M, K), bqb.view(N, K), sab, sbb, out=out), flush)
        tm = tmq = float("nan")
        if K % 128 == 0:
            amx, sfa = quantize_mxfp8(ab)
            bmx, sfb = quantize_mxfp8(bb)
            tm = timeit(lambda: gemm_mxfp8(amx, bmx, sfa, sfb, out=out), flush)
            tmq = timeit(lambda: quantize_mxfp8(ab), flush)
        tq = timeit(lambda: quantize_fp8_rowwise(ab), flush)
        fl = 2.0 * M * N * K
        row = {"shape": (M, N, K), "fp8_ms": round(t8, 4), "fp8_tflops": round(fl / t8 / 1e9, 1), "fp8_block_ms": round(tb, 4),
               "fp8_block_tflops": round(fl / tb / 1e9, 1), "mxfp8_ms": round(tm, 4), "mxfp8_tflops": round(fl / tm / 1e9, 1),
               "mxfp8_quantize_activation_ms": round(tmq, 4), "bf16_ms": round(t16, 4), "bf16_tflops": round(fl / t16 / 1e9, 1),
               "cublaslt_fp8_ms": round(tl, 4), "cublaslt_fp8_tflops": round(fl / tl / 1e9, 1), "rowwise_quantize_activation_ms": round(tq, 4)}
        rows.append(row)
        print(row, flush=True)
    doc = {"what": "fp8 GEMM variants vs bf16 and cuBLASLt fp8; median of 10, L2 flushed", "rows": rows, "clocks": sampler.stop()}
    print(json.dumps(doc["clocks"]))
    if "--out" in sys.argv:
        with open(sys.argv[sys.argv.index("--out") + 1], "w") as f:
            json.dump(doc, f, indent=1)


if __name__ == "__main__":
    main()
