from .gated_activation import CUDAGatedActivation  # noqa: F401
