"""Time flash attention fwd / bwd (B=2, S=4096, 32 heads, d=128, causal) with CUDA events; prints TFLOP/s."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from deepspeedsyclsupport_b200.ops import flash_attention as FA
B, S, h, d = 2, 4096, 32, 128
qkv = torch.randn(B * S, 3 * h * d, device="cuda", dtype=torch.bfloat16)
out, lse = FA.flash_fwd(qkv, B, S, h, h, d, True)
dout = torch.randn_like(out)
def t(fn, n=10):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(n): fn()
    e.record(); torch.cuda.synchronize()
    return s.elapsed_time(e) / n
f = 4 * B * h * S * S * d / 2
tf = t(lambda: FA.flash_fwd(qkv, B, S, h, h, d, True))
tb = t(lambda: FA.flash_bwd(dout, qkv, out, lse, B, S, h, h, d, True))
print(f"fwd {tf:.3f} ms {f / tf / 1e9:.0f} TFLOP/s | bwd {tb:.3f} ms {2.5 * f / tb / 1e9:.0f} TFLOP/s | env skip_dq={os.environ.get('DSB200_BWD_SKIP_DQ')}")
