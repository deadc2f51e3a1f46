from .cuda_fp_ln_base import CUDAFPLNBase  # noqa: F401
from .cuda_ln import CUDAFPLN  # noqa: F401
from .cuda_post_ln import CUDAFPPostLN  # noqa: F401
from .cuda_pre_ln import CUDAFPPreLN  # noqa: F401
