from .compress import init_compression, redundancy_clean
from .helper import convert_conv1d_to_linear
from .scheduler import compression_scheduler  # noqa: F401
