"""Fused compute+collective paths vs. the NCCL(+separate kernels) composition, with the achieved fraction of roofline.

Launch:  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \
             tools/fused_paths_bench.py --out gpurun_out/fused_paths_N.json
Timing: CUDA events on the launching stream, `--iters` calls after `--warmup`, max over ranks; a device barrier kernel
(symm.barrier) aligns the ranks before every timed call of the peer-memory kernels, exactly as the engine does.
Roofline of a path = max(compute time at the measured GEMM peak, bytes this GPU must receive (or send) / NVLink
unidirectional bandwidth, HBM bytes / measured copy bandwidth); fraction = roofline / measured.
Paths:
  ag_gemm      ZeRO-3 parameter all-gather fused into the consuming GEMM       vs  NCCL all_gather + GEMM
  rs_norm      gradient reduce-scatter fused with 1/N scale + fp32 accumulate + grad-norm sum-of-squares
                                                                                vs  NCCL reduce_scatter + scale/add + norm
  rs_adam      gradient reduce-scatter fused with the partitioned AdamW update vs  NCCL reduce_scatter + fused Adam
  allgather    peer-pull all-gather of updated bf16 shards                      vs  NCCL all_gather
  allreduce    one-shot peer all-reduce (TP inference, small messages)          vs  NCCL all_reduce
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

NVLINK_UNI_GBS = 900.0  # NVLink-5 per direction per GPU (B200_PROFILING.md)


def peaks():
    """(cuBLAS bf16 burst TFLOP/s, copy GB/s) from the driver-written MEASURED_PEAKS.json; documented fallbacks otherwise."""
    gemm_tflops, copy_gbs = 1600.0, 6000.0
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.isfile(path):
        try:
            d = json.load(open(path))
            gemm_tflops = float(d.get("bf16_tflops", gemm_tflops))
            copy_gbs = float(d.get("hbm_gbs", copy_gbs))
        except Exception:
            pass
    return gemm_tflops, copy_gbs


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--iters", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--out", default=None)
    ap.add_argument("--local_rank", type=int, default=-1)
    a = ap.parse_args()
    import torch
    import deepspeedsyclsupport_b200 as ds
    from deepspeedsyclsupport_b200 import comm as dist
    from deepspeedsyclsupport_b200.ops import flat_optim
    from deepspeedsyclsupport_b200.ops.fused_ops import sumsq
    from deepspeedsyclsupport_b200.ops.gemm import gemm
    from deepspeedsyclsupport_b200.parallel.fused import ag_gemm
    from deepspeedsyclsupport_b200.parallel.symm import SymmetricBuffer
    ds.init_distributed()
    rank, W = dist.get_rank(), dist.get_world_size()
    dev = torch.device("cuda", torch.cuda.current_device())
    gemm_peak, copy_bw = peaks()
    results = []

    def timed(fn, pre=None):
        for _ in range(a.warmup):
            if pre:
                pre()
            fn()
        torch.cuda.synchronize()
        dist.barrier()
        ts = []
        for _ in range(a.iters):
            if pre:
                pre()
            s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            s.record()
            fn()
            e.record()
            torch.cuda.synchronize()
            ts.append(s.elapsed_time(e))
        ts.sort()
        t = torch.tensor([ts[len(ts) // 2]], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def report(name, shape, fused_ms, base_ms, compute_ms, link_ms, hbm_ms, note=""):
        roof = max(compute_ms, link_ms, hbm_ms)
        row = dict(path=name, world=W, shape=shape, fused_ms=round(fused_ms, 4), nccl_composition_ms=round(base_ms, 4),
                   speedup=round(base_ms / fused_ms, 3), roofline_ms=round(roof, 4), roofline_fraction=round(roof / fused_ms, 3),
                   bound="compute" if roof == compute_ms else ("nvlink" if roof == link_ms else "hbm"), note=note)
        results.append(row)
        if rank == 0:
            print(json.dumps(row), flush=True)

    # ------------------------------------------------------------------ 1. AG-GEMM (Llama-7B gate_up projection of one layer)
    M, N, K = 8192, 22016, 4096
    unit = N * K
    S = -(-unit // (W * 64)) * 64
    sb = SymmetricBuffer(dist.get_world_group(), S * 2)
    torch.manual_seed(7)
    full_w = torch.randn(unit, device=dev).bfloat16() * 0.02
    pad = torch.zeros(W * S - unit, device=dev, dtype=torch.bfloat16)
    flat = torch.cat([full_w, pad])
    my_shard = sb.view(0, (S, ), torch.bfloat16)
    my_shard.copy_(flat[rank * S:(rank + 1) * S])
    sb.barrier()
    x = torch.randn(M, K, device=dev, dtype=torch.bfloat16)
    w_local = torch.empty(N, K, device=dev, dtype=torch.bfloat16)
    out = torch.empty(M, N, device=dev, dtype=torch.bfloat16)
    flags = torch.zeros(256, device=dev, dtype=torch.int32)
    gathered = torch.empty(W * S, device=dev, dtype=torch.bfloat16)

    def nccl_path():
        dist.all_gather_into_tensor(gathered, my_shard)
        gemm(x, gathered[:unit].view(N, K), out=out)

    def fused_path():
        ag_gemm(x, w_local, out, sb.ptrs, 0, S, flags, N, K)

    base = timed(nccl_path)
    fused = timed(fused_path, pre=sb.barrier)
    ref = x[:64].float() @ full_w.view(N, K).float().t()
    err = (out[:64].float() - ref).abs().max().item() / ref.abs().max().item()
    compute_ms = 2.0 * M * N * K / (gemm_peak * 1e12) * 1e3
    link_ms = (W - 1) / W * unit * 2 / (NVLINK_UNI_GBS * 1e9) * 1e3
    report("ag_gemm", f"X[{M},{K}] x W[{N},{K}]^T, W sharded {W}-way", fused, base, compute_ms, link_ms, 0.0, note=f"rel_err={err:.1e}")
    torch.cuda.synchronize()
    sb.close()

    # ------------------------------------------------------------------ 2./3. gradient reduce-scatter fused with scale+norm / Adam
    n = 202_383_360 // (W * 64) * 64  # one Llama-7B layer of gradients per rank-slice group
    sb = SymmetricBuffer(dist.get_world_group(), W * n * 2)
    g = sb.view(0, (W * n, ), torch.bfloat16)
    g.copy_(torch.randn(W * n, device=dev).bfloat16() * 1e-3)
    sb.barrier()
    acc = torch.zeros(n, device=dev)
    sq = torch.zeros(1, device=dev)
    tmp = torch.empty(n, device=dev, dtype=torch.bfloat16)

    def nccl_rs_norm():
        dist.reduce_scatter_tensor(tmp, g)
        acc.add_(tmp.float(), alpha=1.0 / W)
        sumsq(acc, out=sq)

    def fused_rs_norm():
        sb.reduce_scatter(acc, 0, torch.bfloat16, n, scale=1.0 / W, accumulate=True, sumsq=sq)

    base = timed(nccl_rs_norm)
    fused = timed(fused_rs_norm, pre=sb.barrier)
    link_ms = (W - 1) * n * 2 / (NVLINK_UNI_GBS * 1e9) * 1e3
    hbm_ms = (n * 2 + n * 8) / (copy_bw * 1e9) * 1e3
    report("rs_norm", f"{W}x{n} bf16 grads -> fp32 shard (+= , sumsq)", fused, base, 0.0, link_ms, hbm_ms)

    p32 = torch.randn(n, device=dev)
    m32, v32 = torch.zeros(n, device=dev), torch.zeros(n, device=dev)
    lp = torch.empty(n, device=dev, dtype=torch.bfloat16)

    def nccl_rs_adam():
        dist.reduce_scatter_tensor(tmp, g)
        flat_optim.adam_step(p32, m32, v32, tmp, lr=1e-4, beta1=0.9, beta2=0.95, eps=1e-8, weight_decay=0.1, adamw_mode=True, bias_correction=True,
                             step=1, lp=lp, inv_scale=None, grad_scale=1.0 / W)

    def fused_rs_adam():
        sb.rs_adam(0, p32, m32, v32, lp, n, scale=1.0 / W, lr=1e-4, beta1=0.9, beta2=0.95, eps=1e-8, weight_decay=0.1, step=1)

    base = timed(nccl_rs_adam)
    fused = timed(fused_rs_adam, pre=sb.barrier)
    hbm_ms = (n * 2 + n * 12 + n * 14) / (copy_bw * 1e9) * 1e3
    report("rs_adam", f"{W}x{n} bf16 grads -> AdamW on the fp32 shard", fused, base, 0.0, link_ms, hbm_ms)

    # ------------------------------------------------------------------ 4. all-gather of updated shards
    full = torch.empty(W * n, device=dev, dtype=torch.bfloat16)
    shard_view = sb.view(0, (n, ), torch.bfloat16)

    def nccl_ag():
        dist.all_gather_into_tensor(full, shard_view)

    def fused_ag():
        sb.all_gather(full, 0, n * 2)

    base = timed(nccl_ag)
    fused = timed(fused_ag, pre=sb.barrier)
    report("allgather", f"{W}x{n} bf16", fused, base, 0.0, link_ms, (W * n * 2) / (copy_bw * 1e9) * 1e3)

    # ------------------------------------------------------------------ 5. one-shot all-reduce (TP decode: 64 tokens x 8192)
    for numel in (64 * 8192, 2048 * 8192):
        part = sb.view(0, (numel, ), torch.bfloat16)
        part.copy_(torch.randn(numel, device=dev).bfloat16())
        dst = torch.empty(numel, device=dev, dtype=torch.bfloat16)
        ref_buf = part.clone()

        def nccl_ar():
            dist.all_reduce(ref_buf)

        def fused_ar():
            sb.all_reduce_bf16(dst, 0, numel, scratch_offset=numel * 2)  # auto: two-shot for >= 4 MiB on > 2 GPUs

        base = timed(nccl_ar)
        fused = timed(fused_ar, pre=sb.barrier)
        link = (W - 1) * numel * 2 / (NVLINK_UNI_GBS * 1e9) * 1e3
        report("allreduce", f"{numel} bf16 ({numel * 2 // 1024} KiB)", fused, base, 0.0, link, (W + 1) * numel * 2 / (copy_bw * 1e9) * 1e3)
    torch.cuda.synchronize()
    sb.close()
    if rank == 0 and a.out:
        os.makedirs(os.path.dirname(os.path.abspath(a.out)), exist_ok=True)
        with open(a.out, "w") as f:
            json.dump({"world": W, "gemm_peak_tflops": gemm_peak, "copy_gbs": copy_bw, "nvlink_uni_gbs": NVLINK_UNI_GBS, "results": results}, f, indent=1)


if __name__ == "__main__":
    main()
