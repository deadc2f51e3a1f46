from .partition_parameters import (GatheredParameters, Init, ZeroParamStatus, ZeroParamType, is_zero_param,  # noqa: F401
                                   register_external_parameter, unregister_external_parameter)
from .tiling import TiledLinear, TiledLinearReturnBias  # noqa: F401
from .mics import MiCS_Init, MiCS_Optimizer  # noqa: F401
from .stage3 import DeepSpeedZeroOptimizer_Stage3, estimate_zero3_model_states_mem_needs_all_cold, estimate_zero3_model_states_mem_needs_all_live  # noqa: F401
from .stage_1_and_2 import DeepSpeedZeroOptimizer, estimate_zero2_model_states_mem_needs_all_cold, estimate_zero2_model_states_mem_needs_all_live  # noqa: F401
