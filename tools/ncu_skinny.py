import torch, sys
sys.path.insert(0, '.')
from deepspeedsyclsupport_b200.ops import gemm as G
x = torch.randn(16, 8192, device='cuda').bfloat16(); w = (torch.randn(7168, 8192, device='cuda') * 0.02).bfloat16()
for _ in range(3): G.gemm(x, w)
torch.cuda.synchronize()
