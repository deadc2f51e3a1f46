import os, sys
sys.path.insert(0, ".")
import torch
from deepspeedsyclsupport_b200.ops import flash_attention as FA
B, S, h, d = 2, 4096, 32, 128
qkv = torch.randn(B * S, 3 * h * d, device="cuda", dtype=torch.bfloat16)
for _ in range(3):
    out, lse = FA.flash_fwd(qkv, B, S, h, h, d, True)
torch.cuda.synchronize()
