"""FusedLamb.  Parity: deepspeed/ops/lamb/fused_lamb.py (max_coeff/min_coeff trust-ratio clamp, eps_inside_sqrt=False)."""
import torch

from ..flat_optim import lamb_step


class FusedLamb(torch.optim.Optimizer):

    def __init__(self, params, lr=1e-3, bias_correction=True, betas=(0.9, 0.999), eps=1e-8, eps_inside_sqrt=False,
                 weight_decay=0.0, max_grad_norm=0.0, max_coeff=10.0, min_coeff=0.01, amsgrad=False):
        if amsgrad:
            raise RuntimeError("FusedLamb does not support the AMSGrad variant.")
        defaults = dict(lr=lr, bias_correction=bias_correction, betas=betas, eps=eps, weight_decay=weight_decay,
                        max_grad_norm=max_grad_norm, max_coeff=max_coeff, min_coeff=min_coeff)
        super().__init__(params, defaults)
        self.eps_mode = 0 if eps_inside_sqrt else 1
        self.lamb_coeffs = []

    @torch.no_grad()
    def step(self, closure=None, grads=None, output_params=None, scale=1.0, grad_norms=None):
        loss = closure() if closure is not None else None
        del self.lamb_coeffs[:]
        for group in self.param_groups:
            beta1, beta2 = group["betas"]
            for p in group["params"]:
                if p.grad is None:
                    continue
                st = self.state[p]
                if len(st) == 0:
                    st["step"] = 0
                    st["exp_avg"] = torch.zeros_like(p.data, dtype=torch.float32)
                    st["exp_avg_sq"] = torch.zeros_like(p.data, dtype=torch.float32)
                st["step"] += 1
                combined_scale = scale
                if group["max_grad_norm"] > 0 and grad_norms is not None:
                    pass
                inv = None
                g = p.grad.data.contiguous().view(-1)
                if combined_scale != 1.0:
                    g = g.float() / combined_scale
                p32 = p.data if p.dtype == torch.float32 else p.data.float()
                coeff = lamb_step(p32.view(-1), st["exp_avg"].view(-1), st["exp_avg_sq"].view(-1), g, lr=group["lr"], beta1=beta1,
                                  beta2=beta2, eps=group["eps"], weight_decay=group["weight_decay"],
                                  bias_correction=group["bias_correction"], step=st["step"], max_coeff=group["max_coeff"],
                                  min_coeff=group["min_coeff"], inv_scale=inv)
                self.lamb_coeffs.append(coeff)
                if p32 is not p.data:
                    p.data.copy_(p32)
        return loss

    def get_lamb_coeffs(self):
        return [float(c) for c in self.lamb_coeffs]
