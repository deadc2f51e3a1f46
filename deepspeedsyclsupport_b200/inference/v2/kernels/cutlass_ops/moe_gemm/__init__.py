from .moe_gemm import MoEGEMM  # noqa: F401
from .mixed_moe_gemm import MixedMoEGEMM  # noqa: F401
