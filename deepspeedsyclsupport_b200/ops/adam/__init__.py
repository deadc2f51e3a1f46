from .cpu_adam import DeepSpeedCPUAdam  # noqa: F401
from .fused_adam import FusedAdam  # noqa: F401
