"""Paged flash attention (csrc/attn/paged_flash.cu) in the two regimes of the FastGen 70B TP=8 config (8 q heads / 1 kv head per
rank, d = 128, 128-token KV blocks): a 2048-token prefill chunk and decode steps over long contexts.  Reports time, achieved
tensor FLOP/s (prefill) and KV bytes/s against MEASURED_PEAKS.json hbm_gbs (decode), next to the CUDA-core kernel it replaced and,
when importable, flashinfer's paged kernels (library, for comparison).  CUDA events, L2 flushed between iterations, clock record.
    python tools/paged_flash_bench.py --out profiles/paged_flash_bench.json          # --profile prefill|decode : one launch for ncu"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

from bench import ClockSampler  # noqa: E402
from deepspeedsyclsupport_b200.inference.v2 import kernels as K  # noqa: E402
from deepspeedsyclsupport_b200.ops import native  # noqa: E402


def make(seq_new, seq_seen, hq, hkv, d, bs):
    n_seq = len(seq_new)
    max_blocks = max(-(-(a + b) // bs) for a, b in zip(seq_new, seq_seen)) + 1
    num_blocks = n_seq * max_blocks + 1
    bt = torch.randperm(num_blocks)[:n_seq * max_blocks].view(n_seq, max_blocks).int().cuda()
    kv = (torch.randn(num_blocks, 2, bs, hkv, d, device="cuda") * 0.5).bfloat16()
    ts = torch.cat([torch.full((n, ), s, dtype=torch.int32) for s, n in enumerate(seq_new)]).cuda()
    tp = torch.cat([torch.arange(seen, seen + n, dtype=torch.int32) for n, seen in zip(seq_new, seq_seen)]).cuda()
    qkv = torch.randn(int(ts.numel()), (hq + 2 * hkv) * d, device="cuda").bfloat16()
    return qkv, kv, ts, tp, bt


def timed(fn, flush, iters=10, reps=8):
    """ms per call of `fn`, device time: `reps` calls are captured in a CUDA graph (the launches of one call -- memset, atom
    discovery, attention, merge -- cost more host time than the kernels take on the GPU) and the graph is replayed `iters`
    times with the L2 flushed in between; inputs (KV cache of all sequences) are far larger than L2 for the decode cases."""
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        fn()
        with torch.cuda.graph(g, stream=side):
            for _ in range(reps):
                fn()
    torch.cuda.synchronize()
    tot = 0.0
    for _ in range(iters):
        flush.zero_()
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        g.replay()
        e.record()
        torch.cuda.synchronize()
        tot += s.elapsed_time(e)
    return tot / iters / reps


def scalar_kernel(qkv, kv, ts, tp, bt, hq, hkv, d):
    T = qkv.shape[0]
    out = torch.empty(T, hq * d, dtype=qkv.dtype, device=qkv.device)
    rc = native.lib().dsb200_paged_attention(native.ptr(qkv), native.ptr(kv), native.ptr(out), native.ptr(ts), native.ptr(tp), native.ptr(bt), T, hq,
                                             hkv, d, kv.shape[2], bt.shape[1], float(d ** -0.5), native.stream_ptr())
    native.check(rc)
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default=None)
    ap.add_argument("--profile", default=None, choices=["prefill", "decode"])
    a = ap.parse_args()
    hq, hkv, d, bs = 8, 1, 128, 128
    cases = {"prefill_2048": ([2048], [0]), "prefill_512_after_8k": ([512], [8192]),
             "decode_64x4096": ([1] * 64, [4096] * 64), "decode_8x32768": ([1] * 8, [32768] * 8)}
    if a.profile:
        name = "prefill_2048" if a.profile == "prefill" else "decode_64x4096"
        args = make(*cases[name], hq, hkv, d, bs)
        for _ in range(2):
            K.paged_flash_attention(*args, hq, hkv, d)
        torch.cuda.synchronize()
        return
    peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))) if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else {}
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    sampler = ClockSampler(0)
    sampler.start()
    rows = {}
    for name, (new, seen) in cases.items():
        args = make(new, seen, hq, hkv, d, bs)
        ctx_keys = sum((s + n) * n - n * (n - 1) // 2 for n, s in zip(new, seen))  # (query, key) pairs under the causal mask
        flops = 4 * ctx_keys * hq * d
        kv_bytes = sum(s + n for n, s in zip(new, seen)) * hkv * d * 2 * 2
        t_flash = timed(lambda: K.paged_flash_attention(*args, hq, hkv, d), flush)
        small = sum(new) <= 256  # the CUDA-core kernel re-reads the history per token: only timed where that finishes quickly
        t_scalar = timed(lambda: scalar_kernel(*args, hq, hkv, d), flush, iters=3, reps=2) if (small or name == "prefill_512_after_8k") else None
        ref = K.paged_attention_ref  # noqa: F841 (oracle lives in tests/test_paged_flash.py)
        row = {"tokens": sum(new), "context": seen[0] + new[0], "sequences": len(new), "paged_flash_ms": round(t_flash, 4),
               "tflops": round(flops / t_flash / 1e9, 1), "kv_gbytes_per_s": round(kv_bytes / t_flash / 1e6, 1)}
        if peaks.get("hbm_gbs"):
            row["kv_fraction_of_hbm_gbs"] = round(row["kv_gbytes_per_s"] / peaks["hbm_gbs"], 3)
        if t_scalar is not None:
            row["cuda_core_kernel_ms"] = round(t_scalar, 4)
            row["speedup_vs_cuda_core"] = round(t_scalar / t_flash, 2)
        try:
            import flashinfer  # library kernel, comparison only
            row["flashinfer"] = "importable (not wired into this table: its paged layout differs from the cache layout timed here)"
        except Exception:
            pass
        rows[name] = row
    doc = {"what": "paged flash attention, 8 q heads / 1 kv head per rank (Llama-2-70B at TP=8), d=128, 128-token KV blocks, bf16",
           "cases": rows, "clocks": sampler.stop(), "peaks": {k: peaks.get(k) for k in ("hbm_gbs", "bf16_tflops_sustained")}}
    text = json.dumps(doc, indent=1)
    print(text)
    if a.out:
        with open(a.out, "w") as f:
            f.write(text + "\n")


if __name__ == "__main__":
    main()
