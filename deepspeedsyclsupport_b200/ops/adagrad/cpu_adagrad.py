"""DeepSpeedCPUAdagrad.  Parity: deepspeed/ops/adagrad/cpu_adagrad.py."""
import torch

from ..flat_optim import adagrad_step


class DeepSpeedCPUAdagrad(torch.optim.Optimizer):
    optimizer_id = 0

    def __init__(self, model_params, lr=1e-2, eps=1e-10, weight_decay=0, amsgrad=False, fp32_optimizer_states=True):
        super().__init__(model_params, dict(lr=lr, eps=eps, weight_decay=weight_decay, amsgrad=amsgrad))
        self.opt_id = DeepSpeedCPUAdagrad.optimizer_id
        DeepSpeedCPUAdagrad.optimizer_id += 1

    @torch.no_grad()
    def step(self, closure=None, fp16_param_groups=None):
        loss = closure() if closure is not None else None
        if fp16_param_groups is not None and fp16_param_groups and type(fp16_param_groups[0]) is not list:
            fp16_param_groups = [fp16_param_groups]
        for gi, group in enumerate(self.param_groups):
            for pi, p in enumerate(group["params"]):
                if p.grad is None:
                    continue
                st = self.state[p]
                if len(st) == 0:
                    st["step"] = 0
                    st["exp_avg_sq"] = torch.zeros_like(p.data, dtype=torch.float32)
                st["step"] += 1
                p32 = p.data if p.dtype == torch.float32 else p.data.float()
                adagrad_step(p32.view(-1), st["exp_avg_sq"].view(-1), p.grad.data.contiguous().view(-1), lr=group["lr"],
                             eps=group["eps"], weight_decay=group["weight_decay"])
                if p32 is not p.data:
                    p.data.copy_(p32)
                if fp16_param_groups is not None:
                    fp16_param_groups[gi][pi].data.copy_(p.data)
        return loss
