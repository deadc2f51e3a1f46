"""(parity: deepspeed/module_inject/containers/features/split_qkv.py)  Architectures whose q / k / v are separate modules: the
hybrid engine needs to know the parts so LoRA adapters on them can be fused into the fused QKV."""
from abc import abstractmethod

import torch

from .hybrid_engine import HybridEngineContainer


class HybridSplitQKVContainer(HybridEngineContainer):

    def set_attention(self, qkvw, qkvb, dense_w, dense_b):
        super().set_attention(qkvw, qkvb, dense_w, dense_b)
        self.set_q_k_v()

    @abstractmethod
    def set_q_k_v(self):
        """Set self.qw, self.qb, self.kw, self.kb, self.vw, self.vb from the client layer."""

    def attention_qkv_mp(self, mp_replace, reversed_dim=False):
        if getattr(self, "qw", None) is None:
            return super().attention_qkv_mp(mp_replace, reversed_dim=reversed_dim)
        self.qkvw = torch.cat([self.qw, self.kw, self.vw], dim=0)
        if self.qb is not None:
            self.qkvb = torch.cat([self.qb, self.kb, self.vb], dim=0)
        super().attention_qkv_mp(mp_replace, reversed_dim=reversed_dim)
