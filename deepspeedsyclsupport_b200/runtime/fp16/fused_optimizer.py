"""FP16_Optimizer: fp16 parameters, flat fp32 master weights, static / dynamic loss scaling with overflow skip.
Parity: deepspeed/runtime/fp16/fused_optimizer.py:24-480.  Stage-0 configuration of the unified flat store; overflow
detection, unscale, clip and the Adam update run on device without a host round-trip except the skip decision."""
import torch

from ..zero.config import DeepSpeedZeroConfig
from ..zero.sharded_optimizer import ZeroShardedOptimizer


class FP16_Optimizer(ZeroShardedOptimizer):

    def __init__(self, init_optimizer, deepspeed=None, static_loss_scale=1.0, dynamic_loss_scale=False, initial_dynamic_scale=2**32,
                 dynamic_loss_args=None, verbose=True, mpu=None, clip_grad=0.0, fused_adam_legacy=False, has_moe_layers=False, timers=None,
                 module=None, dp_process_group=None):
        module = module if module is not None else getattr(deepspeed, "module", None)
        if module is None:
            raise ValueError("FP16_Optimizer needs `module=` or `deepspeed=` (an engine) to build the flat store")
        if dynamic_loss_scale and dynamic_loss_args is None:
            dynamic_loss_args = {"init_scale": initial_dynamic_scale, "scale_window": 1000, "min_scale": 1, "delayed_shift": 1}
        super().__init__(module, init_optimizer, stage=0, zero_config=DeepSpeedZeroConfig(stage=0), dtype=torch.float16, dp_group=dp_process_group,
                         mpu=mpu, clip_grad=clip_grad, timers=timers, static_loss_scale=static_loss_scale, dynamic_loss_scale=dynamic_loss_scale,
                         dynamic_loss_args=dynamic_loss_args)
        self.deepspeed = deepspeed
        self.fused_adam_legacy = fused_adam_legacy
        self.has_moe_layers = has_moe_layers
        self.verbose = verbose

    @property
    def cur_scale(self):
        return self.loss_scale

    @property
    def fp16_groups(self):
        return [[p for u in self.units for p in u.params]]

    def unscale_and_clip_grads(self, *a, **k):
        pass  # folded into the fused update (device-side coefficient)

    def refresh_fp32_params(self):
        self.refresh_from_params()
