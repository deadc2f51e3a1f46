"""(parity: deepspeed/module_inject/containers/features/megatron.py)  Megatron-LM v2 checkpoints store the fused QKV per head as
(q, k, v) triples; the blocks want [Q; K; V]."""
from abc import ABC

from ...policy import _transpose


class MegatronContainer(ABC):

    def __init__(self, **kwargs):
        super().__init__(**kwargs)
        self.megatron_v2 = self.policy.is_megatron_v2

    def _align_qkv_transposed(self, x):
        return _transpose(x, heads=self.num_attention_heads)

    def transpose(self):
        super().transpose()
        if self.megatron_v2:
            self.qkvw = self._align_qkv_transposed(self.qkvw.data)
            if self.qkvb is not None:
                self.qkvb = self._align_qkv_transposed(self.qkvb.data)
