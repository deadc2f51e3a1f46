from .mixed_gemm import MixedGEMM  # noqa: F401
