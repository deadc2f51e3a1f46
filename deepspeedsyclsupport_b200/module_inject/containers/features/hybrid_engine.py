"""(parity: deepspeed/module_inject/containers/features/hybrid_engine.py)  Hybrid (RLHF) engine support: the inference block
shares storage with the training layer, LoRA deltas are fused before generation and unfused after."""
from abc import ABC
from typing import List, Tuple

import torch


class HybridEngineContainer(ABC):

    def initialize_tensors(self, enable_training=False):
        super().initialize_tensors(enable_training=enable_training)
        self.set_lora_params()

    def set_lora_params(self):
        """[(weight, [lora_right, lora_left, scaling])] for every projection that carries a LoRA adapter."""
        self.lora_params = []

    def get_lora_matched_pair(self) -> List[Tuple[torch.Tensor, list]]:
        return [(w, l) for w, l in getattr(self, "lora_params", []) if len(l) == 3]

    def _delta(self, lora):
        right, left, scaling = lora
        return scaling * torch.matmul(left.t(), right.t())

    def fuse_lora(self):
        for weight, lora in self.get_lora_matched_pair():
            weight.data += self._delta(lora).to(weight.dtype)

    def unfuse_lora(self):
        for weight, lora in self.get_lora_matched_pair():
            weight.data -= self._delta(lora).to(weight.dtype)

    def apply_tensor_parallelism(self, mp_replace, reversed_dim=False, **kwargs):
        super().apply_tensor_parallelism(mp_replace, **kwargs)

    def release_memory(self):
        """Drop the container's references to the client tensors (the block keeps its own copies / views)."""
        self.qkvw = self.qkvb = self.dense_w = self.dense_b = None
        self._h4h_w = self._h4h_b = self._4hh_w = self._4hh_b = None

    def set_params_wo_copy(self, Z3_enabled=False):
        """Point the block's parameters at the training tensors instead of copying (single-GPU hybrid engine)."""
        a, m = self.module.attention, self.module.mlp
        for owner, name, src in ((a, "attn_ow", self.dense_w), (a, "attn_ob", self.dense_b), (m, "output_w", self._4hh_w),
                                 (m, "output_b", self._4hh_b), (m, "attn_nw", self.attn_nw), (m, "attn_nb", self.attn_nb),
                                 (self.module, "norm_w", self.input_nw), (self.module, "norm_b", self.input_nb)):
            if src is not None and getattr(owner, name) is not None and tuple(getattr(owner, name).shape) == tuple(src.shape):
                getattr(owner, name).data = src.data
