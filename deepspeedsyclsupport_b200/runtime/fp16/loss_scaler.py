"""Loss scaling for fp16 training: a fixed scale, or a dynamic one driven by the overflow flag the ZeRO / fp16 optimizers
compute on the device.

Behavioural parity with the reference's scaler (deepspeed/runtime/fp16/loss_scaler.py:67-230): the scale shrinks by
`scale_factor` on an overflow once the hysteresis budget (`delayed_shift`) is used up, never below `min_scale`; it grows by
`scale_factor` after every `scale_window` overflow-free iterations; `consecutive_hysteresis` refills the budget after any
clean iteration instead of only at growth time.  Attribute names are the reference's because checkpoints and user code
read them.
"""

import torch

from ...utils.logging import logger

# ds_config -> constructor keyword names (runtime/config.py builds `dynamic_loss_args` with these)
INITIAL_LOSS_SCALE = "init_scale"
SCALE_WINDOW = "scale_window"
DELAYED_SHIFT = "delayed_shift"
CONSECUTIVE_HYSTERESIS = "consecutive_hysteresis"
MIN_LOSS_SCALE = "min_scale"


def to_python_float(t):
    return t.item() if hasattr(t, "item") else t[0]


def _finite(x):
    """True when the tensor holds no inf / nan (one reduction, one host read)."""
    total = x.float().sum()
    return bool(torch.isfinite(total).item())


class LossScalerBase:
    """Common surface: `loss_scale`, `backward(loss)`, `update_scale(overflow)`."""
    dynamic = False

    def __init__(self, cur_scale):
        self.cur_scale = cur_scale

    loss_scale = property(lambda self: self.cur_scale)

    def backward(self, loss, retain_graph=False):
        (loss * self.cur_scale).backward(retain_graph=retain_graph)

    def scale_gradient(self, module, grad_in, grad_out):
        s = self.cur_scale
        return tuple(g * s for g in grad_in)

    def update_scale(self, overflow):
        return None


class LossScaler(LossScalerBase):
    """Constant scale; never reports an overflow."""

    def __init__(self, scale=1):
        LossScalerBase.__init__(self, scale)

    def has_overflow(self, params):
        return False

    _has_inf_or_nan = staticmethod(lambda x: False)


class DynamicLossScaler(LossScalerBase):
    dynamic = True

    def __init__(self, init_scale=2 ** 32, scale_factor=2.0, scale_window=1000, min_scale=1, delayed_shift=1,
                 consecutive_hysteresis=False, raise_error_at_min_scale=True, dtype=torch.half):
        LossScalerBase.__init__(self, init_scale)
        self.scale_factor, self.scale_window, self.min_scale = scale_factor, scale_window, min_scale
        self.delayed_shift = self.cur_hysteresis = delayed_shift
        self.consecutive_hysteresis = consecutive_hysteresis
        self.raise_error_at_min_scale = raise_error_at_min_scale
        self.dtype = dtype
        self.cur_iter, self.last_overflow_iter = 0, -1

    @staticmethod
    def _has_inf_or_nan(x):
        try:
            return not _finite(x)
        except RuntimeError as err:  # the sum itself overflowed the conversion
            if "value cannot be converted" in str(err):
                return True
            raise

    def has_overflow_serial(self, params):
        return any(p.grad is not None and self._has_inf_or_nan(p.grad.data) for p in params)

    # ---- the state machine
    def _say(self, text):
        if self.dtype == torch.half:
            from ... import comm as dist
            logger.info(f"[deepspeed] OVERFLOW! Rank {dist.get_rank()} Skipping step. Attempted loss scale: {int(self.cur_scale)}, {text}")

    def _on_overflow(self):
        budget_left = self.delayed_shift != 1 and self.cur_hysteresis != 1
        if budget_left:
            self._say(f"but hysteresis is {self.cur_hysteresis}. Reducing hysteresis to {self.cur_hysteresis - 1}")
            self.cur_hysteresis -= 1
        else:
            if self.raise_error_at_min_scale and self.cur_scale == self.min_scale:
                raise Exception("Current loss scale already at minimum - cannot decrease scale anymore. Exiting run.")
            lowered = max(self.min_scale, self.cur_scale / self.scale_factor)
            self._say(f"reducing to {int(lowered)}")
            self.cur_scale = lowered
        self.last_overflow_iter = self.cur_iter

    def _on_clean(self):
        window_full = (self.cur_iter - self.last_overflow_iter) % self.scale_window == 0
        if self.consecutive_hysteresis or window_full:
            self.cur_hysteresis = self.delayed_shift
        if window_full:
            self.cur_scale *= self.scale_factor

    def update_scale(self, overflow):
        (self._on_overflow if overflow else self._on_clean)()
        self.cur_iter += 1


def CreateLossScaler(dtype, static_loss_scale, dynamic_scaling, dynamic_loss_args):
    """fp16 + dynamic -> DynamicLossScaler(**dynamic_loss_args); anything else -> a constant scale (1.0 unless fp16)."""
    if dtype != torch.half:
        return LossScaler(scale=1.0)
    if not dynamic_scaling:
        return LossScaler(scale=static_loss_scale)
    return DynamicLossScaler(dtype=dtype, **(dynamic_loss_args or {}))
