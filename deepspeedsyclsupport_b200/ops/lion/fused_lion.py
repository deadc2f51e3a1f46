"""FusedLion on the flat Lion kernel (csrc/optim/fused_optim.cu).  Parity: deepspeed/ops/lion/fused_lion.py."""
import torch

from ..flat_optim import lion_step


class FusedLion(torch.optim.Optimizer):

    def __init__(self, params, lr=1e-3, betas=(0.9, 0.999), weight_decay=0.0, set_grad_none=True):
        super().__init__(params, dict(lr=lr, betas=betas, weight_decay=weight_decay))
        self.set_grad_none = set_grad_none

    def zero_grad(self, set_to_none=None):
        super().zero_grad(set_to_none=self.set_grad_none if set_to_none is None else set_to_none)

    @torch.no_grad()
    def step(self, closure=None):
        loss = closure() if closure is not None else None
        for group in self.param_groups:
            beta1, beta2 = group["betas"]
            for p in group["params"]:
                if p.grad is None:
                    continue
                st = self.state[p]
                if len(st) == 0:
                    st["exp_avg"] = torch.zeros_like(p.data, dtype=torch.float32)
                p32 = p.data if p.dtype == torch.float32 else p.data.float()
                lion_step(p32.view(-1), st["exp_avg"].view(-1), p.grad.data.contiguous().view(-1), lr=group["lr"], beta1=beta1,
                          beta2=beta2, weight_decay=group["weight_decay"])
                if p32 is not p.data:
                    p.data.copy_(p32)
        return loss
