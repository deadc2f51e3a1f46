"""Per-phase clock stamps of one CTA of the flash backward (debug hook dsb200_flash_fwd_trace)."""
import sys
sys.path.insert(0, ".")
import torch
from deepspeedsyclsupport_b200.ops import flash_attention as FA, native
B, S, h, d = 2, 4096, 32, 128
qkv = torch.randn(B * S, 3 * h * d, device="cuda", dtype=torch.bfloat16)
out, lse = FA.flash_fwd(qkv, B, S, h, h, d, True)
dout = torch.randn_like(out)
FA.flash_bwd(dout, qkv, out, lse, B, S, h, h, d, True)
buf = torch.zeros(8192, dtype=torch.int64, device="cuda")
native.lib().dsb200_flash_fwd_trace(buf.data_ptr())
FA.flash_bwd(dout, qkv, out, lse, B, S, h, h, d, True)
torch.cuda.synchronize()
native.lib().dsb200_flash_fwd_trace(None)
t = buf.view(-1, 16).cpu()
t0 = int(t[0, 0])
names = ["mma:start", "mma:SdP_issued", "mma:PdS_seen", "mma:all_issued", "sm:SdP_seen", "sm:PdS_arrive", "sm:mma2_seen", "sm:dq_drained", "sm:dq_tma", "sm:stage_free", "sm:pre_fence"]
print(" ".join(n.rjust(14) for n in ["step"] + names))
for step in range(10, 16):
    print(" ".join(str(x).rjust(14) for x in [step] + [int(t[step, e]) - t0 for e in range(11)]))
print("per-step period:", (int(t[15, 0]) - int(t[10, 0])) / 5)
