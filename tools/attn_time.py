"""Head-to-head of the repo's flash attention (csrc/attn) against torch SDPA backends (flash = FA-2, cuDNN) on B200.

Causal, bf16, d=128; shapes: the flagship training shape (B2 / S4096 / 32 heads) and a long-context one (B1 / S32768 / 8 heads).
CUDA-event timing after warm-up, L2 flushed between iterations, `nvidia-smi` clock/throttle record for the whole run
(bench.ClockSampler).  Prints one JSON document (also written to --out).
    python tools/attn_time.py --out profiles/attn_head_to_head.json
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
import torch.nn.functional as F  # noqa: E402
from torch.nn.attention import SDPBackend, sdpa_kernel  # noqa: E402

from bench import ClockSampler  # noqa: E402
from deepspeedsyclsupport_b200.ops import flash_attention as FA  # noqa: E402


def timed(fn, flush, iters=10, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    total = 0.0
    for _ in range(iters):
        flush.zero_()  # 256 MiB write: evicts the 126 MB L2
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        fn()
        e.record()
        torch.cuda.synchronize()
        total += s.elapsed_time(e)
    return total / iters


def run_shape(B, S, h, d, flush):
    flops_fwd = 4 * B * h * S * S * d / 2  # causal: half of the S x S tiles
    qkv = torch.randn(B * S, 3 * h * d, device="cuda", dtype=torch.bfloat16)
    out, lse = FA.flash_fwd(qkv, B, S, h, h, d, True)
    dout = torch.randn_like(out)
    rows = {}
    t_f = timed(lambda: FA.flash_fwd(qkv, B, S, h, h, d, True), flush)
    t_b = timed(lambda: FA.flash_bwd(dout, qkv, out, lse, B, S, h, h, d, True), flush)
    rows["own (csrc/attn, tcgen05)"] = {"fwd_ms": t_f, "bwd_ms": t_b}
    q, k, v = (x.reshape(B, S, h, d).transpose(1, 2).contiguous().requires_grad_(True) for x in qkv.view(B * S, 3, h * d).unbind(1))
    g = dout.view(B, S, h, d).transpose(1, 2).contiguous()
    for name, backend in (("torch SDPA flash (FA-2)", SDPBackend.FLASH_ATTENTION), ("torch SDPA cuDNN", SDPBackend.CUDNN_ATTENTION)):
        try:
            with sdpa_kernel(backend):
                o = F.scaled_dot_product_attention(q, k, v, is_causal=True)
                ref_err = (o.transpose(1, 2).reshape(B * S, h * d).float() - out.float()).abs().max().item()
                tf = timed(lambda: F.scaled_dot_product_attention(q, k, v, is_causal=True), flush)

                def fb():
                    oo = F.scaled_dot_product_attention(q, k, v, is_causal=True)
                    oo.backward(g)
                    q.grad = k.grad = v.grad = None

                tfb = timed(fb, flush)
            rows[name] = {"fwd_ms": tf, "bwd_ms": tfb - tf, "max_abs_diff_vs_own_fwd": ref_err}
        except Exception as ex:  # backend not built / shape unsupported
            rows[name] = {"unavailable": f"{type(ex).__name__}: {str(ex)[:120]}"}
    for r in rows.values():
        if "fwd_ms" in r:
            r["fwd_tflops"] = round(flops_fwd / r["fwd_ms"] / 1e9, 1)
            r["bwd_tflops"] = round(2.5 * flops_fwd / r["bwd_ms"] / 1e9, 1)
            r["fwd_ms"], r["bwd_ms"] = round(r["fwd_ms"], 4), round(r["bwd_ms"], 4)
    return {"shape": {"B": B, "S": S, "heads": h, "d": d, "causal": True, "dtype": "bf16"}, "results": rows}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default=None)
    a = ap.parse_args()
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    sampler = ClockSampler(0)
    sampler.start()
    doc = {"what": "flash attention fwd / bwd, own kernels vs torch SDPA backends; CUDA events, L2 flushed between iterations",
           "shapes": [run_shape(2, 4096, 32, 128, flush), run_shape(1, 32768, 8, 128, flush)]}
    doc["clocks"] = sampler.stop()
    text = json.dumps(doc, indent=1)
    print(text)
    if a.out:
        os.makedirs(os.path.dirname(os.path.abspath(a.out)), exist_ok=True)
        with open(a.out, "w") as f:
            f.write(text + "\n")


if __name__ == "__main__":
    main()
