"""Target for `ncu -k regex:flash_|gemm_grouped`: one forward + backward of the attention kernels on the flagship shape and one
grouped expert GEMM (Mixtral gate_up, 4096 tokens)."""
import sys
sys.path.insert(0, ".")
import torch
from deepspeedsyclsupport_b200.ops import flash_attention as FA
from deepspeedsyclsupport_b200.ops.gemm import moe_gemm
B, S, h, d = 2, 4096, 32, 128
qkv = torch.randn(B * S, 3 * h * d, device="cuda", dtype=torch.bfloat16)
for _ in range(2):
    out, lse = FA.flash_fwd(qkv, B, S, h, h, d, True)
    dout = torch.randn_like(out)
    FA.flash_bwd(dout, qkv, out, lse, B, S, h, h, d, True)
E, K, N, R = 8, 4096, 28672, 8192
w = torch.randn(E, N, K, device="cuda", dtype=torch.bfloat16) * 0.02
x = torch.randn(R, K, device="cuda", dtype=torch.bfloat16)
cum = torch.full((E, ), R // E, device="cuda").cumsum(0)
for _ in range(2):
    moe_gemm(x, w, cum)
torch.cuda.synchronize()
