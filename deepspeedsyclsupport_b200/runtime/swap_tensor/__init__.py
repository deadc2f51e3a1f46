from .optimizer_swapper import OptimizerStateSwapper  # noqa: F401
from .utils import SwapBufferPool, swap_in_tensors, swap_out_tensors  # noqa: F401
from .aio_config import get_aio_config  # noqa: F401,E402
from .partitioned_optimizer_swapper import PartitionedOptimizerSwapper, PipelinedOptimizerSwapper  # noqa: F401,E402
from .partitioned_param_swapper import AsyncPartitionedParameterSwapper, PartitionedParamStatus  # noqa: F401,E402
