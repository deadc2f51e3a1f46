"""DeepSpeedCPULion: Lion for host-resident (ZeRO-Offload) parameters, optionally copying the updated fp32 values into 16-bit
device parameters.  Parity: deepspeed/ops/lion/cpu_lion.py."""
import torch

from .fused_lion import FusedLion


class DeepSpeedCPULion(FusedLion):
    optimizer_id = 0

    def __init__(self, model_params, lr=1e-3, betas=(0.9, 0.999), weight_decay=0, fp32_optimizer_states=True):
        super().__init__(model_params, lr=lr, betas=betas, weight_decay=weight_decay)
        self.opt_id = DeepSpeedCPULion.optimizer_id
        DeepSpeedCPULion.optimizer_id += 1

    @torch.no_grad()
    def step(self, closure=None, fp16_param_groups=None):
        loss = super().step(closure)
        if fp16_param_groups is not None:
            if fp16_param_groups and type(fp16_param_groups[0]) is not list:
                fp16_param_groups = [fp16_param_groups]
            for gi, group in enumerate(self.param_groups):
                for pi, p in enumerate(group["params"]):
                    fp16_param_groups[gi][pi].data.copy_(p.data)
        return loss
