from .bias_activation import CUDABiasActivation  # noqa: F401
