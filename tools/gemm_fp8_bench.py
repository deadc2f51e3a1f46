"""fp8 row-scaled GEMM (dsb200_gemm_fp8) vs the bf16 CTA-pair GEMM and torch._scaled_mm (cuBLASLt fp8) on the training shapes.
CUDA events, L2 flushed between iterations."""
import sys

import torch

sys.path.insert(0, ".")
from deepspeedsyclsupport_b200.ops.gemm import gemm, gemm_fp8, quantize_fp8_rowwise  # noqa: E402


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
        fl = 2.0 * M * N * K
        print({"shape": (M, N, K), "fp8_ms": round(t8, 4), "fp8_tflops": round(fl / t8 / 1e9, 1), "bf16_ms": round(t16, 4),
               "bf16_tflops": round(fl / t16 / 1e9, 1), "cublaslt_fp8_ms": round(tl, 4), "cublaslt_fp8_tflops": round(fl / tl / 1e9, 1)}, flush=True)


if __name__ == "__main__":
    main()
