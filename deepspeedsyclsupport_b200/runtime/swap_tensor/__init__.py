from .optimizer_swapper import OptimizerStateSwapper  # noqa: F401
from .utils import SwapBufferPool, swap_in_tensors, swap_out_tensors  # noqa: F401
