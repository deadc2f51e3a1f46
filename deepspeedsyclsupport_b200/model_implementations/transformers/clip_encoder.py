"""`DSClipEncoder`: the CLIP text encoder of a diffusion pipeline behind CUDA graphs -- one per call pattern (the pipeline
alternates conditional / unconditional prompts) (reference: deepspeed/model_implementations/transformers/clip_encoder.py)."""
import torch

from ..features.cuda_graph import CUDAGraph


class DSClipEncoder(CUDAGraph, torch.nn.Module):

    def __init__(self, enc, enable_cuda_graph=False):
        torch.nn.Module.__init__(self)
        CUDAGraph.__init__(self, enable_cuda_graph=enable_cuda_graph)
        if hasattr(enc, "text_model") and hasattr(enc.text_model, "_build_causal_attention_mask"):
            enc.text_model._build_causal_attention_mask = self._build_causal_attention_mask
        self.enc = enc
        self.device = getattr(enc, "device", None)
        self.dtype = getattr(enc, "dtype", None)
        self.config = getattr(enc, "config", None)

    @staticmethod
    def _build_causal_attention_mask(bsz, seq_len, dtype):
        # built on the current device so that it can live inside a captured graph
        dev = "cuda" if torch.cuda.is_available() else "cpu"
        mask = torch.full((bsz, seq_len, seq_len), torch.finfo(dtype).min, dtype=dtype, device=dev)
        return mask.triu_(1).unsqueeze(1)

    def _forward(self, *inputs, **kwargs):
        return self.enc(*inputs, **kwargs)

    def forward(self, *inputs, **kwargs):
        return self._graphed_call(self._forward, *inputs, **kwargs)
