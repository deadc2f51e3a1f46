"""Tiny driver for ncu: runs the tcgen05 GEMM (fwd / dgrad / wgrad layouts) and flash attention a few times."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from deepspeedsyclsupport_b200.ops.gemm import gemm
from deepspeedsyclsupport_b200.ops import flash_attention as FA

M, N, K = 8192, 4096, 4096
a = torch.randn(M, K, device="cuda", dtype=torch.bfloat16)
b = torch.randn(N, K, device="cuda", dtype=torch.bfloat16)
c = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
for _ in range(4):
    gemm(a, b, out=c)
B, S, h, d = 2, 4096, 32, 128
qkv = torch.randn(B * S, 3 * h * d, device="cuda", dtype=torch.bfloat16)
for _ in range(3):
    out, lse = FA.flash_fwd(qkv, B, S, h, h, d, True)
dout = torch.randn_like(out)
for _ in range(3):
    FA.flash_bwd(dout, qkv, out, lse, B, S, h, h, d, True)
torch.cuda.synchronize()
print("ok")
