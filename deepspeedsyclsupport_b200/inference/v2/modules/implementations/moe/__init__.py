from .cutlass_multi_gemm import DSMultiGemmMoE
