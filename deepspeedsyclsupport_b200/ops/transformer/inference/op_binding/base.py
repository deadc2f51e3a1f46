"""(parity: deepspeed/ops/transformer/inference/op_binding/base.py)  The reference binds each op to one of several compiled
functions by dtype; here every op dispatches at call time: bf16 CUDA tensors take the tcgen05 GEMM / fused kernels, everything
else the torch path with identical math."""
import torch

from ..config import DeepSpeedInferenceConfig


class BaseOp(torch.nn.Module):
    inference_module = None

    def __init__(self, config: DeepSpeedInferenceConfig):
        super().__init__()
        self.config = config


def linear_nt(x, w, b=None):
    """x [..., in] @ w[out, in]^T (+ b)"""
    from ..ds_attention import _linear
    return _linear(x, w, b)
