from .comm import *  # noqa: F401,F403
from .comm import _get_world_group  # noqa: F401
from .reduce_op import ReduceOp  # noqa: F401
