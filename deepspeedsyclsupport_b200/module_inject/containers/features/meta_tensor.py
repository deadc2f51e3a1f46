"""(parity: deepspeed/module_inject/containers/features/meta_tensor.py)  Containers that can be filled from a checkpoint state
dict without materialising the client model (it lives on the meta device)."""
from abc import ABC, abstractmethod


class MetaTensorContainer(ABC):

    def __init__(self, **kwargs):
        super().__init__(**kwargs)
        self.is_meta = False
        self.ckpt_load_enabled = True

    def initialize_tensors(self, enable_training=False):
        super().initialize_tensors(enable_training=enable_training)
        self.is_meta = self.qkvw is not None and self.qkvw.is_meta

    def apply_tensor_parallelism(self, mp_replace, **kwargs):
        if self.is_meta:
            return  # nothing to copy yet: load_params fills the block from the checkpoint
        super().apply_tensor_parallelism(mp_replace, **kwargs)

    def copy_data_to_new_module(self):
        if self.is_meta:
            return
        super().copy_data_to_new_module()

    def transpose(self):
        if self.is_meta:
            return
        super().transpose()

    @abstractmethod
    def load_params(self, module, sd, weight_quantizer, mp_replace, prefix):
        """Copy this layer's tensors from `sd` (keys start with `prefix`) into `module`."""
