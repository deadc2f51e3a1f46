"""1-bit Adam: dense Adam for `freeze_step` warm-up steps, then the variance is frozen and only the momentum is
communicated, 1 bit per element with error feedback.  Parity: deepspeed/runtime/fp16/onebit/adam.py:14-306."""
import types

import torch

from .... import comm as dist
from .base import CompressedCommOptimizer


class OnebitAdam(CompressedCommOptimizer):

    def __init__(self, params, deepspeed=None, lr=1e-3, freeze_step=100000, bias_correction=True, betas=(0.9, 0.999), eps=1e-8,
                 eps_inside_sqrt=False, weight_decay=0., max_grad_norm=0., amsgrad=False, cuda_aware=False, comm_backend_name="nccl"):
        if amsgrad:
            raise RuntimeError("1-bit Adam does not support the AMSGrad variant.")
        defaults = dict(lr=lr, bias_correction=bias_correction, betas=betas, eps=eps, weight_decay=weight_decay, max_grad_norm=max_grad_norm)
        super().__init__(params, defaults)
        self.eps_mode = 0 if eps_inside_sqrt else 1
        self.comm_time = 0.0
        self.step_time = 0.0
        self.ave_step = 1
        self.bk_time = 0.0
        self.adam_freeze_key = False
        self.initialize = False
        self.freeze_step = freeze_step
        self._init_comm(deepspeed, comm_backend_name, cuda_aware)

    @torch.no_grad()
    def step(self, closure=None, grads=None):
        loss = closure() if closure is not None else None
        if grads is None:
            grads_group = [None] * len(self.param_groups)
        elif isinstance(grads, types.GeneratorType) or type(grads[0]) != list:
            grads_group = [list(grads)]
        else:
            grads_group = grads
        last_step = 0
        for group, grads_this_group in zip(self.param_groups, grads_group):
            beta1, beta2 = group["betas"]
            for p, grad in self._grads(group, grads_this_group):
                state = self.state[p]
                if len(state) == 0:
                    state["step"] = 0
                    state["exp_avg"] = torch.zeros_like(p.data)
                    state["exp_avg_sq"] = torch.zeros_like(p.data)
                if not self.initialize or (self.adam_freeze_key and "worker_error" not in state):
                    self._new_error_buffers(state, p)
                    self.adam_freeze_key = True
                exp_avg, exp_avg_sq = state["exp_avg"], state["exp_avg_sq"]
                state["step"] += 1
                last_step = state["step"]
                if self.adam_freeze_key is False:
                    exp_avg.mul_(beta1).add_(grad, alpha=1 - beta1)
                    exp_avg_sq.mul_(beta2).addcmul_(grad, grad, value=1 - beta2)
                elif group.get("non_freeze", False):
                    dist.all_reduce(grad)
                    grad.mul_(1 / dist.get_world_size())
                    exp_avg.mul_(beta1).add_(grad, alpha=1 - beta1)
                    exp_avg_sq.mul_(beta2).addcmul_(grad, grad, value=1 - beta2)
                else:
                    if self.initialize:
                        exp_avg.mul_(beta1).add_(grad, alpha=1 - beta1)
                    if self.size > 1:
                        self.comm_backend_handle.compressed_allreduce(exp_avg, state["worker_error"], state["server_error"], self._local_rank())
                    self._apply_mask(group, exp_avg)
                if self.initialize:
                    update = exp_avg / (exp_avg_sq.sqrt() + group["eps"])
                    if group["weight_decay"] > 0.0:
                        update += group["weight_decay"] * p.data
                    p.add_(update, alpha=-group["lr"])
            if not self.initialize:
                for p in group["params"]:
                    self.state[p].pop("worker_error", None)
                    self.state[p].pop("server_error", None)
        if not self.initialize:
            # the very first call only exercises the compressed path once (buffer allocation / NCCL warm-up)
            self.adam_freeze_key = False
            self.initialize = True
            return loss
        if self.adam_freeze_key is False and last_step >= self.freeze_step:
            self.adam_freeze_key = True
            self._set_dense_allreduce(False)
        return loss

    def load_state_dict(self, state_dict):
        self._keep_user_masks(state_dict)
        super().load_state_dict(state_dict)
        first = self.state[self.param_groups[0]["params"][0]]
        warm = first.get("step", 0) < self.freeze_step
        if warm and self.adam_freeze_key:
            self.adam_freeze_key = False
            self._set_dense_allreduce(True)
        elif not warm and not self.adam_freeze_key:
            self.adam_freeze_key = True
            self._set_dense_allreduce(False)
        self._drop_error_buffers()
