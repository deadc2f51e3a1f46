"""diffusers UNet2DConditionModel -> DSUNet (CUDA-graph replay wrapper) (parity: deepspeed/module_inject/containers/unet.py)."""
import torch

from ..policy import DSPolicy


class UNetPolicy(DSPolicy):

    def __init__(self):
        super().__init__()
        try:
            import diffusers
            self._orig_layer_class = diffusers.models.unet_2d_condition.UNet2DConditionModel
        except (ImportError, AttributeError):
            self._orig_layer_class = None

    def match(self, module):
        return self._orig_layer_class is not None and isinstance(module, self._orig_layer_class)

    def match_replaced(self, module):
        from ...model_implementations.diffusers.unet import DSUNet
        return isinstance(module, DSUNet)

    def apply(self, module, enable_cuda_graph=True):
        from ...model_implementations.diffusers.unet import DSUNet
        return DSUNet(module, enable_cuda_graph=enable_cuda_graph)

    def attention(self, client_module):
        """(qkv weight or separate q / k / v, output weight, output bias, fused?) of a diffusers CrossAttention module."""
        qw, kw, vw = client_module.to_q.weight, client_module.to_k.weight, client_module.to_v.weight
        ow, ob = client_module.to_out[0].weight, client_module.to_out[0].bias
        if qw.shape[1] == kw.shape[1]:  # self-attention: one fused projection
            return torch.nn.Parameter(torch.cat((qw, kw, vw), dim=0), requires_grad=False), ow, ob, True
        return qw, kw, vw, ow, ob, False  # cross-attention: keys / values come from the encoder width
