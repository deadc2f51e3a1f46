"""Diagnostics: weight-streaming GEMM time vs the number of K slices (DSB200_SKINNY_SPLITS), decode shapes of Llama-2-70B at TP = 8."""
import sys
sys.path.insert(0, '.')
import torch
from deepspeedsyclsupport_b200.ops import gemm as G
flush = torch.empty(256 << 20, dtype=torch.uint8, device='cuda')
def timed(fn, iters=15):
    for _ in range(3): fn()
    torch.cuda.synchronize(); tot = 0.0
    for _ in range(iters):
        flush.zero_()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record(); torch.cuda.synchronize(); tot += a.elapsed_time(b)
    return tot / iters
for name, M, N, K, cands in [("gate_up", 16, 7168, 8192, (1, 2, 3, 4, 5, 8)), ("down", 16, 8192, 3584, (1, 2, 3, 4, 6, 7)), ("qkv", 16, 1280, 8192, (4, 8, 14, 16, 29)),
                             ("o", 16, 8192, 1024, (1, 2, 4)), ("gate_up x32", 32, 7168, 8192, (2, 5))]:
    x = torch.randn(M, K, device='cuda').bfloat16(); w = (torch.randn(N, K, device='cuda') * 0.02).bfloat16()
    res = {}
    for s in cands:
        G._FORCE_SPLITS = s
        res[s] = round(timed(lambda: G.gemm(x, w)) * 1e3, 1)
    G._FORCE_SPLITS = 0
    print(name, "tiles", -(-N // 128), "us by splits:", res, flush=True)
