"""Per-phase clock stamps of one CTA of the flash forward (debug hook dsb200_flash_fwd_trace): prints the softmax / MMA hand-off
latencies per key block for the heaviest CTA."""
import sys
sys.path.insert(0, ".")
import torch
from deepspeedsyclsupport_b200.ops import flash_attention as FA, native
B, S, h, d = 2, 4096, 32, 128
qkv = torch.randn(B * S, 3 * h * d, device="cuda", dtype=torch.bfloat16)
FA.flash_fwd(qkv, B, S, h, h, d, True)
buf = torch.zeros(8192, dtype=torch.int64, device="cuda")
native.lib().dsb200_flash_fwd_trace(buf.data_ptr())
FA.flash_fwd(qkv, B, S, h, h, d, True)
torch.cuda.synchronize()
native.lib().dsb200_flash_fwd_trace(None)
t = buf.view(-1, 2, 8).cpu()
t0 = int(t[0, 0, 0])
print("j  tile | S_seen ld_done max_done P_arrive | mma:P_seen PV_issued | softmax X | wait_for_S")
prev = {0: None, 1: None}
for j in range(8, 16):
    for tile in (0, 1):
        e = [int(x) - t0 for x in t[j, tile]]
        X = e[3] - e[0]
        w = e[0] - prev[tile] if prev[tile] is not None else 0
        prev[tile] = e[3]
        print(j, tile, "|", e[0], e[1] - e[0], e[2] - e[1], e[3] - e[2], "|", e[4] - e[3], e[5] - e[4], "|", X, "|", w)
