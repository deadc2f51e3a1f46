"""`DSVAE`: encode / decode / forward of the diffusion VAE, each behind its own CUDA graph
(reference: deepspeed/model_implementations/diffusers/vae.py)."""
import torch

from ..features.cuda_graph import CUDAGraph


class DSVAE(CUDAGraph, torch.nn.Module):

    def __init__(self, vae, enable_cuda_graph=True):
        torch.nn.Module.__init__(self)
        CUDAGraph.__init__(self, enable_cuda_graph=enable_cuda_graph)
        self.vae = vae
        self.config = getattr(vae, "config", None)
        self.device = getattr(vae, "device", None)
        self.dtype = getattr(vae, "dtype", None)
        self.vae.requires_grad_(requires_grad=False)

    def _decode(self, x, return_dict=True, generator=None):
        return self.vae.decode(x, return_dict=return_dict)

    def _encode(self, x, return_dict=True):
        return self.vae.encode(x, return_dict=return_dict)

    def _forward(self, sample, timestamp=None, encoder_hidden_states=None, return_dict=True):
        return self.vae(sample, timestamp, encoder_hidden_states, return_dict)

    def decode(self, *inputs, **kwargs):
        return self._graphed_call(self._decode, *inputs, **kwargs)

    def encode(self, *inputs, **kwargs):
        return self._graphed_call(self._encode, *inputs, **kwargs)

    def forward(self, *inputs, **kwargs):
        return self._graphed_call(self._forward, *inputs, **kwargs)
