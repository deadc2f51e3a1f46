"""1-bit LAMB: LAMB warm-up, then compressed momentum all-reduce with frozen + adaptively rescaled trust ratios.
Parity: deepspeed/runtime/fp16/onebit/lamb.py:15-443.

One structural difference from the reference: all momenta are views into ONE flat buffer from the start (the reference
re-links them lazily with _flatten_dense_tensors at the first compressed step), so the compressed all-reduce always runs
once over the fused buffer."""
import types

import numpy as np
import torch

from .base import CompressedCommOptimizer


class OnebitLamb(CompressedCommOptimizer):

    def __init__(self, params, deepspeed=None, lr=1e-3, freeze_step=100000, bias_correction=True, betas=(0.9, 0.999), eps=1e-8,
                 eps_inside_sqrt=False, weight_decay=0., max_grad_norm=0., max_coeff=10.0, min_coeff=0.01, amsgrad=False,
                 cuda_aware=False, comm_backend_name="nccl", coeff_beta=0.9, factor_max=4.0, factor_min=0.5, factor_threshold=0.1):
        if amsgrad:
            raise RuntimeError("1-bit Lamb does not support the AMSGrad variant.")
        defaults = dict(lr=lr, bias_correction=bias_correction, betas=betas, eps=eps, weight_decay=weight_decay, max_grad_norm=max_grad_norm,
                        max_coeff=max_coeff, min_coeff=min_coeff)
        super().__init__(params, defaults)
        self.eps_mode = 0 if eps_inside_sqrt else 1
        self.lamb_freeze_key = False
        self.initialize = False
        self.freeze_step = freeze_step
        self.coeff_beta = coeff_beta
        self.factor_max = factor_max
        self.factor_min = factor_min
        self.factor_threshold = factor_threshold
        self._init_comm(deepspeed, comm_backend_name, cuda_aware)
        self.exp_avg_flat = None
        self.worker_error = None
        self.server_error = None
        self.lamb_coeffs = []

    def _all_params(self):
        for group in self.param_groups:
            for p in group["params"]:
                yield group, p

    def _ensure_state(self):
        """Momenta live as views of one flat (padded) buffer."""
        if self.exp_avg_flat is not None:
            return
        ps = [p for _, p in self._all_params()]
        total = sum(p.numel() for p in ps)
        self.corrected_tensor_size = self._corrected_size(total)
        self.server_chunk_size = self.corrected_tensor_size // self.size
        dev = ps[0].device
        self.exp_avg_flat = torch.zeros(self.corrected_tensor_size, dtype=torch.float32, device=dev)
        off = 0
        for p in ps:
            state = self.state[p]
            prev = state.get("exp_avg")
            view = self.exp_avg_flat[off:off + p.numel()].view_as(p.data)
            if prev is not None:
                view.copy_(prev)
            state["exp_avg"] = view
            state.setdefault("step", 0)
            state.setdefault("lamb_coeff_freeze", 0.0)
            state.setdefault("last_factor", 1.0)
            if "exp_avg_sq" not in state:
                state["exp_avg_sq"] = torch.zeros_like(p.data, dtype=torch.float32)
                state["exp_avg_sq_fresh"] = torch.zeros_like(p.data, dtype=torch.float32)
            off += p.numel()

    def _error_buffers(self):
        if self.worker_error is None:
            dev = self.exp_avg_flat.device
            self.worker_error = torch.zeros(self.corrected_tensor_size, device=dev)
            self.server_error = torch.zeros(self.server_chunk_size, device=dev)

    @torch.no_grad()
    def step(self, closure=None, grads=None):
        loss = closure() if closure is not None else None
        if grads is None:
            grads_group = [None] * len(self.param_groups)
        elif isinstance(grads, types.GeneratorType) or type(grads[0]) != list:
            grads_group = [list(grads)]
        else:
            grads_group = grads
        self._ensure_state()
        del self.lamb_coeffs[:]
        if not self.initialize:
            self.lamb_freeze_key = True
        exp_avg_last_step = None
        if self.lamb_freeze_key:
            exp_avg_last_step = {p: self.state[p]["exp_avg"].detach().clone() for _, p in self._all_params()}
            first = next(self._all_params())[1]
            if "scaling_coeff" not in self.state[first]:
                # equalise the momentum magnitudes across tensors so one shared 1-bit scale fits all of them
                scales = {p: (torch.linalg.norm(self.state[p]["exp_avg"]) / np.sqrt(p.numel())).item() for _, p in self._all_params()}
                united = sum(scales.values()) / len(scales)
                for _, p in self._all_params():
                    self.state[p]["scaling_coeff"] = united / scales[p] if scales[p] != 0 else 1.0
        last_step = 0
        for group, grads_this_group in zip(self.param_groups, grads_group):
            beta1, beta2 = group["betas"]
            max_coeff, min_coeff = group["max_coeff"], group["min_coeff"]
            for p, grad in self._grads(group, grads_this_group):
                state = self.state[p]
                exp_avg, exp_avg_sq, exp_avg_sq_fresh = state["exp_avg"], state["exp_avg_sq"], state["exp_avg_sq_fresh"]
                state["step"] += 1
                last_step = state["step"]
                if self.lamb_freeze_key is False:
                    exp_avg.mul_(beta1).add_(grad, alpha=1 - beta1)
                    exp_avg_sq.mul_(beta2).addcmul_(grad, grad, value=1 - beta2)
                    if state["step"] == self.freeze_step:
                        exp_avg_sq_fresh.copy_(exp_avg_sq)
                    if self.initialize:
                        weight_norm = p.data.pow(2).sum().sqrt()
                        update = exp_avg / (exp_avg_sq.sqrt() + group["eps"])
                        if group["weight_decay"] > 0.0:
                            update += group["weight_decay"] * p.data
                        update_norm = update.pow(2).sum().sqrt()
                        lamb_coeff = 1.0
                        if weight_norm != 0 and update_norm != 0:
                            lamb_coeff = min(max_coeff, max(min_coeff, (weight_norm / update_norm).item()))
                        if lamb_coeff != 1.0:
                            state["lamb_coeff_freeze"] = self.coeff_beta * state["lamb_coeff_freeze"] + (1 - self.coeff_beta) * lamb_coeff
                        self.lamb_coeffs.append(lamb_coeff)
                        p.add_(update, alpha=-group["lr"] * lamb_coeff)
                elif self.initialize:
                    exp_avg.mul_(beta1).add_(grad, alpha=1 - beta1)
                    exp_avg.mul_(state["scaling_coeff"])
        if self.lamb_freeze_key and self.size > 1:
            self._error_buffers()
            self.comm_backend_handle.compressed_allreduce(self.exp_avg_flat, self.worker_error, self.server_error, self._local_rank())
            if not self.initialize:
                self.worker_error = self.server_error = None
        if self.lamb_freeze_key and self.initialize:
            for group in self.param_groups:
                beta1, beta2 = group["betas"]
                for p in group["params"]:
                    state = self.state[p]
                    exp_avg, exp_avg_sq, exp_avg_sq_fresh = state["exp_avg"], state["exp_avg_sq"], state["exp_avg_sq_fresh"]
                    exp_avg.div_(state["scaling_coeff"])
                    self._apply_mask(group, exp_avg)
                    grad_reconstruct = (exp_avg - exp_avg_last_step[p] * beta1) / (1 - beta1)
                    exp_avg_sq_fresh.mul_(beta2).addcmul_(grad_reconstruct, grad_reconstruct, value=1 - beta2)
                    denom = exp_avg_sq.sqrt() + group["eps"]
                    update_prelim = exp_avg / denom
                    update = update_prelim + group["weight_decay"] * p.data if group["weight_decay"] > 0.0 else update_prelim
                    update_norm = update.pow(2).sum().sqrt()
                    denom_real = exp_avg_sq_fresh.sqrt() + group["eps"]
                    factor = (denom / denom_real).max().item()
                    if group["weight_decay"] > 0.0:
                        update_ratio = min(1.0, (update_prelim.pow(2).sum().sqrt() / update_norm).item())
                        factor = factor * update_ratio + (1.0 - update_ratio)
                    factor = min(self.factor_max, max(self.factor_min, factor))
                    factor = min(factor, state["last_factor"] * (1.0 + self.factor_threshold))
                    factor = max(factor, state["last_factor"] * (1.0 - self.factor_threshold))
                    state["last_factor"] = factor
                    lamb_coeff = state["lamb_coeff_freeze"] * factor
                    self.lamb_coeffs.append(lamb_coeff)
                    p.add_(update, alpha=-group["lr"] * lamb_coeff)
        if not self.initialize:
            self.lamb_freeze_key = False
            self.initialize = True
            for _, p in self._all_params():
                self.state[p].pop("scaling_coeff", None)
            return loss
        if self.lamb_freeze_key is False and last_step >= self.freeze_step:
            self.lamb_freeze_key = True
            self._set_dense_allreduce(False)
        return loss

    def load_state_dict(self, state_dict):
        self._keep_user_masks(state_dict)
        super().load_state_dict(state_dict)
        self.exp_avg_flat = None  # loading replaced the momentum tensors: re-link them into a flat buffer lazily
        self.worker_error = self.server_error = None
        first = self.state[self.param_groups[0]["params"][0]]
        if first.get("step", 0) < self.freeze_step:
            if self.lamb_freeze_key is True:
                self.lamb_freeze_key = False
                self._set_dense_allreduce(True)
            for _, p in self._all_params():
                self.state[p]["lamb_coeff_freeze"] = 0.0
                self.state[p]["last_factor"] = 1.0
                self.state[p].pop("scaling_coeff", None)
        else:
            if self.lamb_freeze_key is False:
                self.lamb_freeze_key = True
                self._set_dense_allreduce(False)

    def get_lamb_coeffs(self):
        return self.lamb_coeffs
