"""FP16_UnfusedOptimizer: the per-parameter (unflattened) fp16 wrapper used with LAMB-type optimizers that need
per-tensor norms.  Parity: deepspeed/runtime/fp16/unfused_optimizer.py:22-420.  Here the flat store already keeps
per-parameter boundaries (runs / unit offsets), and FusedLamb computes its per-tensor trust ratios from them, so this is
the same machinery under the reference's class name."""
from .fused_optimizer import FP16_Optimizer


class FP16_UnfusedOptimizer(FP16_Optimizer):

    def __init__(self, init_optimizer, deepspeed=None, static_loss_scale=1.0, dynamic_loss_scale=False, dynamic_loss_args=None, verbose=True,
                 mpu=None, clip_grad=0.0, fused_lamb_legacy=False, **kw):
        super().__init__(init_optimizer, deepspeed=deepspeed, static_loss_scale=static_loss_scale, dynamic_loss_scale=dynamic_loss_scale,
                         dynamic_loss_args=dynamic_loss_args, verbose=verbose, mpu=mpu, clip_grad=clip_grad, **kw)
        self.fused_lamb_legacy = fused_lamb_legacy
