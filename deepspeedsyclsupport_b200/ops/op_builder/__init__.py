from .builder import build_all, build_device, build_host, LIB_DIR  # noqa: F401
from .all_ops import *  # noqa: F401,F403
from .all_ops import ALL_OPS, BUILDERS, OpBuilder  # noqa: F401
