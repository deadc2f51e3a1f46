"""`DSUNet`: the diffusion UNet behind a CUDA graph (reference: deepspeed/model_implementations/diffusers/unet.py)."""
import torch

from ..features.cuda_graph import CUDAGraph


class DSUNet(CUDAGraph, torch.nn.Module):

    def __init__(self, unet, enable_cuda_graph=True):
        torch.nn.Module.__init__(self)
        CUDAGraph.__init__(self, enable_cuda_graph=enable_cuda_graph)
        self.unet = unet
        # attributes the Stable-Diffusion pipeline reads from its unet
        self.in_channels = getattr(unet, "in_channels", getattr(getattr(unet, "config", None), "in_channels", None))
        self.device = getattr(unet, "device", None)
        self.dtype = getattr(unet, "dtype", None)
        self.config = getattr(unet, "config", None)
        self.fwd_count = 0
        self.unet.requires_grad_(requires_grad=False)
        self.unet.to(memory_format=torch.channels_last)

    def _forward(self, sample, timestamp, encoder_hidden_states, return_dict=True, cross_attention_kwargs=None, timestep_cond=None,
                 added_cond_kwargs=None):
        if cross_attention_kwargs:
            return self.unet(sample, timestamp, encoder_hidden_states, return_dict, cross_attention_kwargs=cross_attention_kwargs)
        return self.unet(sample, timestamp, encoder_hidden_states, return_dict)

    def forward(self, *inputs, **kwargs):
        self.fwd_count += 1
        return self._graphed_call(self._forward, *inputs, **kwargs)
