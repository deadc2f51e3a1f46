from .config_v2 import DeepSpeedTPConfig, RaggedInferenceEngineConfig  # noqa: F401
from .engine_factory import build_engine_from_ds_checkpoint, build_engine_from_state_dict, build_hf_engine  # noqa: F401
from .engine_v2 import InferenceEngineV2  # noqa: F401
from .scheduler import SplitFuseScheduler  # noqa: F401
from .scheduling_utils import SchedulingError, SchedulingResult  # noqa: F401
