from .blas_fp_linear import BlasFPLinear, FP8Linear
