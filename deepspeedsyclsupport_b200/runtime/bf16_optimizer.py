"""BF16_Optimizer: bf16 parameters, fp32 master weights + fp32 gradient accumulation, no loss scaling.
Parity: deepspeed/runtime/bf16_optimizer.py:31-460.

In this framework it is the stage-0 configuration of the unified flat store (runtime/zero/sharded_optimizer.py): the
fp32 master / optimizer state live in one flat region, gradients accumulate in fp32, the update is the fused flat
kernel.  The class exists so scripts that construct or isinstance-check BF16_Optimizer keep working."""
import torch

from .. import comm as dist
from .zero.config import DeepSpeedZeroConfig
from .zero.sharded_optimizer import ZeroShardedOptimizer


class BF16_Optimizer(ZeroShardedOptimizer):

    def __init__(self, init_optimizer, param_names=None, mpu=None, clip_grad=0.0, norm_type=2, allgather_bucket_size=5000000000,
                 dp_process_group=None, timers=None, grad_acc_dtype=torch.float32, graph_harvesting=False, module=None, zero_config=None):
        assert norm_type == 2, "only the L2 norm is supported"
        if module is None:
            raise ValueError("BF16_Optimizer needs `module=` (the flat store is built per module unit)")
        super().__init__(module, init_optimizer, stage=0, zero_config=zero_config or DeepSpeedZeroConfig(stage=0), dtype=torch.bfloat16,
                         dp_group=dp_process_group, mpu=mpu, clip_grad=clip_grad, timers=timers,
                         grad_accum_dtype=grad_acc_dtype)
        self.param_names = param_names
        self.graph_harvesting = graph_harvesting

    # names used by the reference engine / user scripts
    def update_hp_grads(self, clear_lp_grads=False):
        self.end_micro_step(True)

    def clear_hp_grads(self):
        self.zero_grad()

    def clear_lp_grads(self):
        for u in self.units:
            for p in u.params:
                p.grad = None

    def update_lp_params(self):
        self._post_step()

    def get_grads_for_norm(self, for_clipping=False):
        return [self.grad_shard]

    @property
    def bf16_groups(self):
        return [[p for u in self.units for p in u.params]]
