"""DeepSpeedCPUAdam -- host Adam for ZeRO-Offload (AVX-512, OpenMP).  Parity: deepspeed/ops/adam/cpu_adam.py:13-181
(constructor, step(closure, fp16_param_groups) semantic: optional 16-bit copies updated in the same pass)."""
import torch

from ..flat_optim import adam_step


class DeepSpeedCPUAdam(torch.optim.Optimizer):
    optimizer_id = 0

    def __init__(self, model_params, lr=1e-3, bias_correction=True, betas=(0.9, 0.999), eps=1e-8, weight_decay=0,
                 amsgrad=False, adamw_mode=True, fp32_optimizer_states=True):
        default_args = dict(lr=lr, betas=betas, eps=eps, weight_decay=weight_decay, bias_correction=bias_correction,
                            amsgrad=amsgrad)
        super().__init__(model_params, default_args)
        self.opt_id = DeepSpeedCPUAdam.optimizer_id
        DeepSpeedCPUAdam.optimizer_id += 1
        self.adam_w_mode = adamw_mode
        self.fp32_optimizer_states = fp32_optimizer_states

    def __setstate__(self, state):
        super().__setstate__(state)
        for group in self.param_groups:
            group.setdefault("amsgrad", False)

    @torch.no_grad()
    def step(self, closure=None, fp16_param_groups=None):
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        device = torch.device("cpu")
        if type(fp16_param_groups) is list and fp16_param_groups and type(fp16_param_groups[0]) is not list:
            fp16_param_groups = [fp16_param_groups]
        for group_id, group in enumerate(self.param_groups):
            for param_id, p in enumerate(group["params"]):
                if p.grad is None:
                    continue
                assert p.device == device, f"CPUAdam param is on {p.device} and must be 'cpu', make sure you enabled 'offload_optimizer': 'cpu' in your ZeRO config."
                state = self.state[p]
                if len(state) == 0:
                    state["step"] = 0
                    state["exp_avg"] = torch.zeros_like(p.data, dtype=torch.float32, device=device)
                    state["exp_avg_sq"] = torch.zeros_like(p.data, dtype=torch.float32, device=device)
                state["step"] += 1
                beta1, beta2 = group["betas"]
                lp = None
                if fp16_param_groups is not None:
                    lp_dev = fp16_param_groups[group_id][param_id]
                    lp = torch.empty(p.numel(), dtype=lp_dev.dtype)
                p32 = p.data if p.dtype == torch.float32 else p.data.float()
                adam_step(p32.view(-1), state["exp_avg"].view(-1), state["exp_avg_sq"].view(-1), p.grad.data.contiguous().view(-1),
                          lr=group["lr"], beta1=beta1, beta2=beta2, eps=group["eps"], weight_decay=group["weight_decay"],
                          adamw_mode=self.adam_w_mode, bias_correction=group["bias_correction"], step=state["step"], lp=lp)
                if p32 is not p.data:
                    p.data.copy_(p32)
                if lp is not None:
                    fp16_param_groups[group_id][param_id].data.copy_(lp.view_as(fp16_param_groups[group_id][param_id]), non_blocking=True)
        return loss
