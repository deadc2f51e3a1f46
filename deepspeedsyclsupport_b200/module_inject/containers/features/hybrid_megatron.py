"""(parity: deepspeed/module_inject/containers/features/hybrid_megatron.py)"""
from .hybrid_engine import HybridEngineContainer
from .megatron import MegatronContainer


class HybridMegatronContainer(MegatronContainer, HybridEngineContainer):

    def _align_qkv(self, x):
        """Megatron-v2 per-head (q, k, v) layout -> [Q; K; V] for the block; the inverse is applied when weights flow back."""
        return self._align_qkv_transposed(x)
