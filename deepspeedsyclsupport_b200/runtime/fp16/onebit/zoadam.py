"""0/1 Adam: variance updates at exponentially growing intervals (dense steps), 1-bit compressed gradient all-reduce in
between, then (after var_freeze_step) local steps with a periodically synchronised momentum accumulator.
Parity: deepspeed/runtime/fp16/onebit/zoadam.py:14-359."""
import types

import torch

from .base import CompressedCommOptimizer


class ZeroOneAdam(CompressedCommOptimizer):

    def __init__(self, params, deepspeed=None, lr=1e-3, bias_correction=True, betas=(0.9, 0.999), eps=1e-8, eps_inside_sqrt=False,
                 weight_decay=0., max_grad_norm=0., var_freeze_step=100000, var_update_scaler=16, local_step_scaler=32678,
                 local_step_clipper=16, amsgrad=False, cuda_aware=False, comm_backend_name="nccl"):
        if amsgrad:
            raise RuntimeError("0/1 Adam does not support the AMSGrad variant.")
        defaults = dict(lr=lr, bias_correction=bias_correction, betas=betas, eps=eps, weight_decay=weight_decay, max_grad_norm=max_grad_norm)
        super().__init__(params, defaults)
        self.eps_mode = 0 if eps_inside_sqrt else 1
        self.initialize = False
        self.var_freeze_step = var_freeze_step
        self.var_update_scaler = var_update_scaler
        self.local_step_scaler = local_step_scaler
        self.local_step_clipper = local_step_clipper
        self.freeze_key = False
        self.reinitial_error_buffer = False
        self._init_comm(deepspeed, comm_backend_name, cuda_aware)

    def _compressed(self, t, state):
        return self.comm_backend_handle.compressed_allreduce(t, state["worker_error"], state["server_error"], self._local_rank())

    @torch.no_grad()
    def step(self, closure=None, grads=None):
        loss = closure() if closure is not None else None
        if grads is None:
            grads_group = [None] * len(self.param_groups)
        elif isinstance(grads, types.GeneratorType) or type(grads[0]) != list:
            grads_group = [list(grads)]
        else:
            grads_group = grads
        for group, grads_this_group in zip(self.param_groups, grads_group):
            beta1, beta2 = group["betas"]
            for p, grad in self._grads(group, grads_this_group):
                state = self.state[p]
                if len(state) == 0:
                    state["step"] = 0
                    state["exp_avg"] = torch.zeros_like(p.data)
                    state["exp_avg_sq"] = torch.zeros_like(p.data)
                if not self.initialize or "worker_error" not in state:
                    state["var_interval"] = 1
                    state["var_counter"] = 0
                    state["local_step_interval"] = 1
                    state["local_step_counter"] = 0
                    state["lrs"] = 0
                    self._new_error_buffers(state, p)
                    state["momentum_accumulator"] = torch.zeros_like(p.data)  # the "u" variable of the 0/1 Adam paper
                exp_avg, exp_avg_sq = state["exp_avg"], state["exp_avg_sq"]
                comm_buffer = state["momentum_accumulator"]
                state["step"] += 1
                if self.initialize:
                    if self.freeze_key is False:
                        if state["step"] % state["var_interval"] == 0:
                            # dense step: the engine all-reduced the gradient
                            exp_avg_sq.mul_(beta2).addcmul_(grad, grad, value=1 - beta2)
                            exp_avg.mul_(beta1).add_(grad, alpha=1 - beta1)
                        else:
                            grad_onebit = grad
                            if self.size > 1:
                                grad_onebit = self._compressed(grad.clone(), state)
                                self._apply_mask(group, grad_onebit)
                            exp_avg.mul_(beta1).add_(grad_onebit, alpha=1 - beta1)
                    else:
                        exp_avg.mul_(beta1).add_(grad, alpha=1 - beta1)
                        state["lrs"] += group["lr"]
                else:
                    if self.size > 1:
                        self._compressed(comm_buffer, state)
                        self._apply_mask(group, comm_buffer)
                if self.initialize:
                    denom = exp_avg_sq.sqrt() + group["eps"]
                    update = exp_avg / denom
                    if group["weight_decay"] > 0.0:
                        update += group["weight_decay"] * p.data
                    p.data.add_(update, alpha=-group["lr"])
                    if self.freeze_key is True:
                        comm_buffer.add_(update, alpha=-group["lr"])
                    if state["step"] % state["local_step_interval"] == 0 and self.freeze_key:
                        # synchronise the accumulated local updates through one compressed all-reduce
                        p.data.add_(comm_buffer, alpha=-1)
                        comm_buffer.mul_(denom)
                        if self.size > 1:
                            self._compressed(comm_buffer, state)
                            self._apply_mask(group, comm_buffer)
                        exp_avg.zero_().add_(comm_buffer / state["lrs"], alpha=-1)
                        p.data.add_(comm_buffer / denom)
                        comm_buffer.zero_()
                        state["lrs"] = 0
                    if self.freeze_key is False:
                        if state["step"] % state["var_interval"] == 0:
                            state["var_counter"] += 1
                            if state["var_counter"] == self.var_update_scaler:
                                state["var_counter"] = 0
                                state["var_interval"] *= 2
                        self._set_dense_allreduce((state["step"] + 1) % state["var_interval"] == 0)
                    else:
                        state["local_step_counter"] += 1
                        if state["local_step_counter"] == self.local_step_scaler:
                            state["local_step_counter"] = 0
                            state["local_step_interval"] = min(self.local_step_clipper, state["local_step_interval"] * 2)
            if not self.initialize:
                self.freeze_key = False
                for p in group["params"]:
                    self.state[p].pop("worker_error", None)
                    self.state[p].pop("server_error", None)
        if not self.initialize:
            self.initialize = True
            return loss
        if self.state[self.param_groups[0]["params"][0]]["step"] > self.var_freeze_step:
            self.freeze_key = True
            self._set_dense_allreduce(False)
        if self.freeze_key is True and self.reinitial_error_buffer is False:
            # the error buffers now track a different quantity (accumulated momentum instead of gradients)
            for group in self.param_groups:
                for p in group["params"]:
                    if "worker_error" in self.state[p]:
                        self.state[p]["worker_error"].zero_()
                        self.state[p]["server_error"].zero_()
            self.reinitial_error_buffer = True
        return loss

    def load_state_dict(self, state_dict):
        self._keep_user_masks(state_dict)
        super().load_state_dict(state_dict)
        first = self.state[self.param_groups[0]["params"][0]]
        if first.get("step", 0) < self.var_freeze_step:
            if self.freeze_key is True:
                self.freeze_key = False
                self._set_dense_allreduce(True)
        else:
            if self.freeze_key is False:
                self.freeze_key = True
                self._set_dense_allreduce(False)
        self.reinitial_error_buffer = False
        self._drop_error_buffers()
