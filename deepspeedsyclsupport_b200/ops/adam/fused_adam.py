"""FusedAdam -- list-of-tensors optimizer API over the multi-tensor CUDA kernel.
Parity: deepspeed/ops/adam/fused_adam.py:18-195 (same constructor arguments; amsgrad unsupported; grads of a group
must share one dtype per launch)."""
import ctypes

import torch

from .. import native
from ..flat_optim import adam_ref


class FusedAdam(torch.optim.Optimizer):

    def __init__(self, params, lr=1e-3, bias_correction=True, betas=(0.9, 0.999), eps=1e-8, adam_w_mode=True,
                 weight_decay=0.0, amsgrad=False, set_grad_none=True):
        if amsgrad:
            raise RuntimeError("FusedAdam does not support the AMSGrad variant.")
        defaults = dict(lr=lr, bias_correction=bias_correction, betas=betas, eps=eps, weight_decay=weight_decay)
        super().__init__(params, defaults)
        self.adam_w_mode = 1 if adam_w_mode else 0
        self.set_grad_none = set_grad_none

    def zero_grad(self, set_to_none=None):
        super().zero_grad(set_to_none=self.set_grad_none if set_to_none is None else set_to_none)

    @torch.no_grad()
    def step(self, closure=None, grads=None, output_params=None, scale=None, grad_norms=None, grad_scaler=None):
        if any(p is not None for p in [grads, output_params, scale, grad_norms]):
            raise RuntimeError("FusedAdam has been updated. Simply initialize it identically to torch.optim.Adam, and call step() with no arguments.")
        loss = closure() if closure is not None else None
        for group in self.param_groups:
            if len(group["params"]) == 0:
                continue
            bias_correction = 1 if group["bias_correction"] else 0
            beta1, beta2 = group["betas"]
            if "step" not in group:
                group["step"] = 0
            group["step"] += 1
            buckets = {}
            for p in group["params"]:
                if p.grad is None:
                    continue
                if p.grad.data.is_sparse:
                    raise RuntimeError("FusedAdam does not support sparse gradients, please consider SparseAdam instead")
                state = self.state[p]
                if len(state) == 0:
                    state["exp_avg"] = torch.zeros_like(p.data, dtype=torch.float32)
                    state["exp_avg_sq"] = torch.zeros_like(p.data, dtype=torch.float32)
                buckets.setdefault((p.is_cuda, p.dtype, p.grad.dtype), []).append(p)
            for (is_cuda, pdt, gdt), plist in buckets.items():
                if is_cuda and pdt == torch.float32:
                    self._launch(plist, gdt, group, bias_correction, beta1, beta2)
                else:
                    # 16-bit or host parameters: fp32 math through the oracle path, cast back
                    for p in plist:
                        st = self.state[p]
                        p32 = p.data.float()
                        adam_ref(p32, st["exp_avg"], st["exp_avg_sq"], p.grad.data, group["lr"], beta1, beta2, group["eps"],
                                 group["weight_decay"], self.adam_w_mode, bias_correction, group["step"])
                        p.data.copy_(p32)
        return loss

    def _launch(self, plist, gdt, group, bias_correction, beta1, beta2):
        n = len(plist)
        arr = ctypes.c_void_p * n
        ps = arr(*[p.data_ptr() for p in plist])
        ms = arr(*[self.state[p]["exp_avg"].data_ptr() for p in plist])
        vs = arr(*[self.state[p]["exp_avg_sq"].data_ptr() for p in plist])
        gs = arr(*[p.grad.data.contiguous().data_ptr() if not p.grad.is_contiguous() else p.grad.data_ptr() for p in plist])
        ns = (ctypes.c_longlong * n)(*[p.numel() for p in plist])
        gt = {torch.float32: 0, torch.bfloat16: 1, torch.float16: 2}[gdt]
        rc = native.lib().dsb200_multi_tensor_adam(n, ctypes.cast(ps, ctypes.c_void_p), ctypes.cast(ms, ctypes.c_void_p),
                                                   ctypes.cast(vs, ctypes.c_void_p), ctypes.cast(gs, ctypes.c_void_p),
                                                   ctypes.cast(ns, ctypes.c_void_p), gt, group["lr"], beta1, beta2, group["eps"],
                                                   group["weight_decay"], self.adam_w_mode, bias_correction, group["step"], None,
                                                   None, native.stream_ptr())
        native.check(rc)
