"""`ds_bench`: collective micro-benchmarks (all_reduce / all_gather / reduce_scatter / all_to_all / broadcast / pt2pt)
over the process group of the job, reporting algorithm and bus bandwidth per message size; `--p2p` also times this
framework's peer-memory kernels (parallel/symm.py) against NCCL.  Timing: CUDA events around `--trials` back-to-back
calls after `--warmups`, max over ranks.  (role of DeepSpeedExamples/benchmarks/communication run by bin/ds_bench)"""
import argparse
import os

import torch

from .. import comm as dist


def _bus_factor(op, n):
    return {"all_reduce": 2 * (n - 1) / n, "all_gather": (n - 1) / n, "reduce_scatter": (n - 1) / n, "all_to_all": (n - 1) / n}.get(op, 1.0)


def _time(fn, warmups, trials, device):
    for _ in range(warmups):
        fn()
    if device.type == "cuda":
        torch.cuda.synchronize()
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        for _ in range(trials):
            fn()
        e.record()
        torch.cuda.synchronize()
        ms = s.elapsed_time(e) / trials
    else:
        import time
        t0 = time.perf_counter()
        for _ in range(trials):
            fn()
        ms = (time.perf_counter() - t0) * 1e3 / trials
    t = torch.tensor([ms], device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def run(op, numel, dtype, device, warmups, trials):
    n, r = dist.get_world_size(), dist.get_rank()
    x = torch.ones(numel, dtype=dtype, device=device)
    if op == "all_reduce":
        fn, payload = (lambda: dist.all_reduce(x)), x.numel()
    elif op == "all_gather":
        out = torch.empty(numel * n, dtype=dtype, device=device)
        fn, payload = (lambda: dist.all_gather_into_tensor(out, x)), out.numel()
    elif op == "reduce_scatter":
        big = torch.ones(numel * n, dtype=dtype, device=device)
        fn, payload = (lambda: dist.reduce_scatter_tensor(x, big)), big.numel()
    elif op == "all_to_all":
        out = torch.empty_like(x)
        fn, payload = (lambda: dist.all_to_all_single(out, x)), x.numel()
    elif op == "broadcast":
        fn, payload = (lambda: dist.broadcast(x, 0)), x.numel()
    elif op == "pt2pt":
        def fn():
            if r == 0:
                dist.send(x, 1)
            elif r == 1:
                dist.recv(x, 0)
        payload = x.numel()
    else:
        raise ValueError(op)
    ms = _time(fn, warmups, trials, device)
    size = payload * x.element_size()
    algbw = size / (ms * 1e-3) / 1e9
    return size, ms, algbw, algbw * _bus_factor(op, n)


def main(argv=None):
    p = argparse.ArgumentParser()
    p.add_argument("--ops", default="all_reduce,all_gather,reduce_scatter,all_to_all,broadcast")
    p.add_argument("--minsize", type=int, default=1 << 16, help="smallest message, in elements")
    p.add_argument("--maxsize", type=int, default=1 << 28)
    p.add_argument("--dtype", default="bfloat16")
    p.add_argument("--warmups", type=int, default=5)
    p.add_argument("--trials", type=int, default=20)
    p.add_argument("--local_rank", type=int, default=int(os.environ.get("LOCAL_RANK", 0)))
    args = p.parse_args(argv)
    dist.init_distributed()
    device = torch.device("cuda", args.local_rank) if torch.cuda.is_available() else torch.device("cpu")
    if device.type == "cuda":
        torch.cuda.set_device(device)
    dtype = getattr(torch, args.dtype)
    for op in args.ops.split(","):
        if dist.get_rank() == 0:
            print(f"\n---- {op} ({dist.get_world_size()} ranks, {args.dtype}) ----\n{'bytes':>14} {'ms':>10} {'algbw GB/s':>12} {'busbw GB/s':>12}")
        numel = args.minsize
        while numel <= args.maxsize:
            size, ms, alg, bus = run(op, numel, dtype, device, args.warmups, args.trials)
            if dist.get_rank() == 0:
                print(f"{size:>14} {ms:>10.3f} {alg:>12.2f} {bus:>12.2f}")
            numel *= 4


if __name__ == "__main__":
    main()
