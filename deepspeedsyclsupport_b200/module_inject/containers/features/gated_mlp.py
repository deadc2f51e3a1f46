"""(parity: deepspeed/module_inject/containers/features/gated_mlp.py)  Gated MLPs keep gate and up projections as separate
modules in the client; the block wants them fused [gate; up]."""
from abc import abstractmethod

import torch

from .hybrid_engine import HybridEngineContainer


class HybridGatedMLPContainer(HybridEngineContainer):

    def set_mlp(self, _h4h_w, _h4h_b, _4hh_w, _4hh_b):
        super().set_mlp(_h4h_w, _h4h_b, _4hh_w, _4hh_b)
        self.set_mlp_gate()

    @abstractmethod
    def set_mlp_gate(self):
        """Set self.inter_up_w / inter_up_b / inter_gate_w / inter_gate_b from the client layer."""

    def mlp_inter_mp(self, mp_replace, reversed_dim=False):
        if getattr(self, "inter_up_w", None) is not None:
            self._h4h_w = torch.cat([self.inter_gate_w, self.inter_up_w], dim=0)
            if self.inter_up_b is not None:
                self._h4h_b = torch.cat([self.inter_gate_b, self.inter_up_b], dim=0)
        super().mlp_inter_mp(mp_replace, reversed_dim=reversed_dim)
