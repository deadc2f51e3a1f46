from .rms_norm_base import CUDARMSNormBase  # noqa: F401
from .rms_norm import CUDARMSNorm  # noqa: F401
from .rms_pre_norm import CUDARMSPreNorm  # noqa: F401
