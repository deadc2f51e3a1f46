"""GPU self-check + timing for the tcgen05 GEMM. Each layout runs in its own subprocess so a fault in
one variant cannot poison the others.  Usage: python tools/gemm_check.py [--bench]"""
import json
import os
import subprocess
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def run_case(a_mn, b_mn, bench):
    import torch
    from deepspeedsyclsupport_b200.ops.gemm import gemm
    torch.manual_seed(0)
    res = []
    shapes = [(128, 256, 64), (128, 256, 256), (256, 512, 128), (384, 768, 512), (1000, 1096, 328), (4096, 4096, 4096),
              (512, 128, 256), (192, 64, 128)]
    for (M, N, K) in shapes:
        if (a_mn and M % 8) or (b_mn and N % 8):
            continue
        a = torch.randn((K, M) if a_mn else (M, K), device="cuda", dtype=torch.bfloat16)
        b = torch.randn((K, N) if b_mn else (N, K), device="cuda", dtype=torch.bfloat16)
        A = a.t() if a_mn else a
        Bt = b if b_mn else b.t()
        ref = A.float() @ Bt.float()
        for od in (torch.bfloat16, torch.float32):
            c = gemm(a, b, a_mn=a_mn, b_mn=b_mn, out_dtype=od)
            torch.cuda.synchronize()
            err = (c.float() - ref).abs().max().item()
            rel = err / ref.abs().max().item()
            res.append(dict(M=M, N=N, K=K, out=str(od), max_err=err, rel=rel))
            print(f"a_mn={a_mn} b_mn={b_mn} {M}x{N}x{K} {od}: max_err={err:.4f} rel={rel:.2e}", flush=True)
        acc = torch.randn(M, N, device="cuda", dtype=torch.float32)
        want = acc + ref
        gemm(a, b, a_mn=a_mn, b_mn=b_mn, out=acc, accumulate=True)
        torch.cuda.synchronize()
        err = (acc - want).abs().max().item()
        print(f"a_mn={a_mn} b_mn={b_mn} {M}x{N}x{K} accumulate-f32: max_err={err:.4f}", flush=True)
    if bench:
        for (M, N, K) in [(8192, 8192, 8192), (8192, 4096, 4096), (8192, 11008, 4096), (8192, 4096, 11008),
                          (4096, 32000, 4096)]:
            if a_mn and b_mn:
                a = torch.randn(K, M, device="cuda", dtype=torch.bfloat16)
                b = torch.randn(K, N, device="cuda", dtype=torch.bfloat16)
            else:
                a = torch.randn(M, K, device="cuda", dtype=torch.bfloat16)
                b = torch.randn((K, N) if b_mn else (N, K), device="cuda", dtype=torch.bfloat16)
            c = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
            flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device="cuda")

            def timeit(fn, iters=10):
                for _ in range(3):
                    fn()
                ts = []
                for _ in range(iters):
                    flush.zero_()
                    s = torch.cuda.Event(enable_timing=True)
                    e = torch.cuda.Event(enable_timing=True)
                    s.record()
                    fn()
                    e.record()
                    torch.cuda.synchronize()
                    ts.append(s.elapsed_time(e))
                ts.sort()
                return ts[len(ts) // 2]

            t_own = timeit(lambda: gemm(a, b, a_mn=a_mn, b_mn=b_mn, out=c))
            A = a.t() if a_mn else a
            Bt = b if b_mn else b.t()
            t_lib = timeit(lambda: torch.matmul(A, Bt, out=c))
            fl = 2.0 * M * N * K
            print(f"BENCH a_mn={a_mn} b_mn={b_mn} {M}x{N}x{K}: own {t_own:.3f} ms = {fl/t_own/1e9:.0f} TFLOP/s | "
                  f"cuBLAS {t_lib:.3f} ms = {fl/t_lib/1e9:.0f} TFLOP/s", flush=True)


if __name__ == "__main__":
    if len(sys.argv) >= 4 and sys.argv[1] == "--case":
        run_case(bool(int(sys.argv[2])), bool(int(sys.argv[3])), "--bench" in sys.argv)
        sys.exit(0)
    rc = 0
    for a_mn, b_mn in [(0, 0), (0, 1), (1, 1), (1, 0)]:
        cmd = [sys.executable, os.path.abspath(__file__), "--case", str(a_mn), str(b_mn)] + (["--bench"] if "--bench" in sys.argv else [])
        try:
            r = subprocess.run(cmd, timeout=420)
            rc |= r.returncode
            print(f"== case a_mn={a_mn} b_mn={b_mn} exit {r.returncode}", flush=True)
        except subprocess.TimeoutExpired:
            print(f"== case a_mn={a_mn} b_mn={b_mn} TIMEOUT", flush=True)
            rc |= 1
    sys.exit(rc)
