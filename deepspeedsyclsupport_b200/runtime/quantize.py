"""MoQ: quantise the 16-bit weights after every optimizer step, lowering the bit width over training (parity:
deepspeed/runtime/quantize.py:14-180).

Every weight matrix that compression marked with `start_bits` / `target_bits` / `q_period` is re-quantised (quantise +
dequantise in place, "fake" quantisation) after each step.  Each time the step counter passes the tensor's period its width
drops by one bit and the period doubles -- times `1 + floor(4 * eigenvalue)` when second-order information is available, so
high-curvature blocks are squeezed more slowly.  Right after a drop the quantised weights are blended with the unquantised ones
(`q_mixed_fp16`), the blend ratio moving to fully quantised at `q_change_ratio` per step."""
import math

import torch

from ..ops.quantizer import ds_quantizer
from ..utils.logging import logger

TWO_D_PARAMS = 6


class Quantizer(object):

    def __init__(self, q_groups=1, q_mixed_fp16=False, q_change_ratio=0.01, q_type=0, q_rounding=0, q_verbose=False, q_eigenvalue=False,
                 use_quantizer_kernel=False, layer_num=0):
        self.q_groups = q_groups
        self.q_mixed_fp16 = q_mixed_fp16
        self.q_change_ratio = q_change_ratio
        self.q_type = q_type  # 'symmetric' | 'asymmetric' (the legacy integer form 0 / 1 is accepted)
        self.q_rounding = q_rounding  # 'nearest' | 'stochastic'
        self.qsteps = 0
        self.quantize_real_ratio = 1.000
        self.q_verbose = q_verbose
        self.q_eigenvalue = q_eigenvalue
        self.use_quantizer_kernel = use_quantizer_kernel
        self.layer_num = layer_num

    # ---- configuration helpers
    @property
    def _symmetric(self):
        return self.q_type in (0, "symmetric")

    @property
    def _stochastic(self):
        return self.q_rounding in (1, "stochastic")

    def any_precision_switch(self):
        """Will some tensor lose a bit at the next step?  (Only then is an eigenvalue computation worth its cost.)"""
        if self.layer_num == 0:
            return True
        result = False
        for index in range(self.layer_num):
            if self.q_start_bits[index] != self.q_target_bits:
                next_step = self.qsteps + (TWO_D_PARAMS * (self.layer_num if self.layer_num != 0 else 1))
                if next_step >= self.q_period[index]:
                    result = True
        return result

    # ---- per-step entry point
    def quantize(self, parameter_group, overflow, eigenvalue_enabled, block_eigenvalue={}):
        if overflow and not eigenvalue_enabled:
            return
        self.step()
        self.update_fp16_ratio()
        for group in parameter_group:
            for p in group:
                if p.dim() > 1 and getattr(p, "start_bits", None):
                    eigenvalue, layer_id = (block_eigenvalue or {}).get(id(p), (None, 0))
                    factor = 1 + math.floor(eigenvalue * 4) if eigenvalue is not None else 1
                    p.data = self.compute_quantization(p, layer_id, factor)

    def step(self):
        self.qsteps += 1

    # ---- quantisers (all return a dequantised tensor of the input's shape and dtype)
    def quantize_highbit(self, inputs, num_bits):
        q_range = 2**num_bits
        shape = inputs.shape
        x = inputs.float().reshape(self.q_groups, -1)
        g_min, g_max = x.amin(dim=-1, keepdim=True), x.amax(dim=-1, keepdim=True)
        noise = torch.empty_like(x).uniform_(-0.5, 0.5) if self._stochastic else 0.0
        if self._symmetric:
            scale = 2 * torch.max(g_min.abs(), g_max.abs()) / q_range
            zero_point = 0.0
            q = (x / scale.clamp_min(1e-12) + noise).round().clamp(-(q_range >> 1), (q_range >> 1) - 1) * scale
        else:
            scale = (g_max - g_min) / q_range
            zero_point = (g_min / scale.clamp_min(1e-12)).round() * scale
            q = ((x - zero_point) / scale.clamp_min(1e-12) + noise).round().clamp(0, q_range - 1) * scale + zero_point
        return q.reshape(shape).to(inputs.dtype)

    def quantize_tenary(self, inputs):
        shape = inputs.shape
        x = inputs.float().reshape(self.q_groups, -1)
        n = x.shape[1]
        threshold = 0.7 * x.norm(p=1, dim=1, keepdim=True) / n
        mask = x.abs() > threshold
        alpha = (mask * x).abs().sum(dim=1, keepdim=True) / mask.sum(dim=1, keepdim=True).clamp_min(1)
        return (alpha * x.sign() * mask).reshape(shape).to(inputs.dtype)

    def quantize_binary(self, inputs):
        shape = inputs.shape
        x = inputs.float().reshape(self.q_groups, -1)
        alpha = x.norm(p=1, dim=1, keepdim=True) / x.shape[1]
        return (alpha * x.sign()).reshape(shape).to(inputs.dtype)

    def mixed_fp16_quantize(self, input, input_q, index):
        if self.q_mixed_fp16 and input.start_bits >= input.target_bits - 1:
            return input * self.quantize_real_ratio + (1 - self.quantize_real_ratio) * input_q
        return input_q

    def compute_quantization(self, input, index=0, factor=1):
        # lower the precision once the tensor's period has elapsed; the next period is twice as long (x factor)
        if input.start_bits != input.target_bits and self.qsteps >= input.q_period:
            self.quantize_real_ratio = 1.0
            input.q_period <<= 1
            input.q_period *= factor
            input.start_bits -= 1
            if self.q_verbose:
                logger.info(f"Quantization settings: current bit-precision = {input.start_bits}, step = {self.qsteps}, "
                            f"quantization period = {input.q_period}, index = {index}")
        assert input.start_bits >= input.target_bits, "Quantization bit is lower than target precision bits!"
        data = input.data
        if self.use_quantizer_kernel:
            if input.start_bits <= 2:
                raise ValueError("Quantization bit is too low, please do it without quantization kernel!")
            input_q = ds_quantizer(data.clone(), self.q_groups, input.start_bits, asym=not self._symmetric, sr=self._stochastic)
        elif input.start_bits >= 3:
            input_q = self.quantize_highbit(data, input.start_bits)
        elif input.start_bits == 2:
            assert self._symmetric, "Quantization type is not symmetric"
            assert not self._stochastic, "Stochastic rounding is not supported for ternary quantization"
            input_q = self.quantize_tenary(data)
        else:
            assert self._symmetric, "Quantization type is not symmetric"
            assert not self._stochastic, "Stochastic rounding is not supported for binary quantization"
            input_q = self.quantize_binary(data)
        if self.q_mixed_fp16 and input.start_bits >= input.target_bits - 1:
            return (self.quantize_real_ratio * data + (1 - self.quantize_real_ratio) * input_q).to(data.dtype)
        return input_q

    def update_fp16_ratio(self):
        if self.q_mixed_fp16:
            self.quantize_real_ratio = max(0.0, self.quantize_real_ratio - self.q_change_ratio)
